// CPU-side checker for the drop-in host headers (mm-admm_b200/host/dropin): runs the driver-layer generators and
// shims and dumps their results as text for tests/test_dropin_host.py.  No CUDA call is made (libmmadmm_b200.so is
// linked only because Mesh.h names its symbols).
#include <fstream>
#include <iostream>
#include "lib/json/single_include/nlohmann/json.hpp"
#include "src/MeshUtils.h"
#include "src/HuangFunctional.h"

static void dump(const char *fn, Eigen::MatrixXd &V, Eigen::MatrixXi &F, std::vector<NodeType> &mask) {
    std::ofstream o(fn);
    o.precision(17);
    o << V.rows() << " " << V.cols() << " " << F.rows() << " " << F.cols() << " " << mask.size() << "\n";
    for (int i = 0; i < V.rows(); i++) { for (int j = 0; j < V.cols(); j++) o << V(i, j) << " "; o << "\n"; }
    for (int i = 0; i < F.rows(); i++) { for (int j = 0; j < F.cols(); j++) o << F(i, j) << " "; o << "\n"; }
    for (size_t i = 0; i < mask.size(); i++) o << (int)mask[i] << "\n";
}

static double circlePhi(double x, double y) { return sqrt((x - 0.5) * (x - 0.5) + (y - 0.5) * (y - 0.5)) - 0.35; }

template <int D> struct Ring : MonitorFunction<D> {   // same expression shape as the reference's MEx1 plug-in
    void operator()(Eigen::Vector<double,D> &x, Eigen::Matrix<double,D,D> &M) override {
        Eigen::Vector<double,D> center(Eigen::Vector<double,D>::Constant(0.5));
        M = Eigen::Matrix<double,D,D>::Identity(M.rows(), M.cols());
        M *= (1 + 20.0 / (1 + 20.0 * ((x - center).squaredNorm())));
    }
};

int main(int argc, char **argv) {
    std::string dir = argv[1];
    {   // box meshes
        std::unordered_map<std::string, double> p;
        p["nx"] = 5; p["ny"] = 5; p["xa"] = 0; p["xb"] = 1; p["ya"] = 0; p["yb"] = 1;
        Eigen::MatrixXd V(6 * 6 + 25, 2); Eigen::MatrixXi F(100, 3); std::vector<NodeType> mask(V.rows());
        utils::generateUniformRectMesh<2>(p, &V, &F, &mask, NodeType::BOUNDARY_FREE);
        dump((dir + "/box2.txt").c_str(), V, F, mask);
        p["nz"] = 3; p["nx"] = 3; p["ny"] = 3; p["za"] = 0; p["zb"] = 1;
        Eigen::MatrixXd V3(4 * 4 * 4 + 27, 3); Eigen::MatrixXi F3(12 * 27, 4); std::vector<NodeType> mask3(V3.rows());
        utils::generateUniformRectMesh<3>(p, &V3, &F3, &mask3, NodeType::BOUNDARY_FIXED);
        dump((dir + "/box3.txt").c_str(), V3, F3, mask3);
        // removeRow == removeRows
        Eigen::MatrixXi A(F), B(F);
        utils::removeRow(A, 7); utils::removeRow(A, 3);
        std::vector<char> drop(F.rows(), 0); drop[7] = drop[3] = 1;
        utils::removeRows(B, drop);
        bool same = A.rows() == B.rows();
        for (int i = 0; same && i < A.rows(); i++) for (int j = 0; j < 3; j++) same = same && A(i, j) == B(i, j);
        std::cout << "removeRow " << (same ? "ok" : "MISMATCH") << "\n";
    }
    {   // level-set disc
        std::function<double(double, double)> phi = circlePhi;
        std::vector<int> n = {12, 12};
        std::vector<std::tuple<double, double>> bb = {{0.0, 1.0}, {0.0, 1.0}};
        Eigen::MatrixXd Vc(1, 2), Vp(1, 2); Eigen::MatrixXi F(1, 3); std::vector<NodeType> mask;
        utils::meshFromLevelSetFun(phi, n, bb, &Vc, &Vp, &F, &mask, NodeType::BOUNDARY_FREE);
        dump((dir + "/disc.txt").c_str(), Vp, F, mask);
    }
    {   // monitor table through the reference's evaluateAtVertices signature
        Eigen::MatrixXd X(3, 2); Eigen::MatrixXi F(1, 3); Eigen::MatrixXd M(3, 4);
        X << 0.1, 0.2, 0.5, 0.5, 0.9, 0.3;
        Ring<2> mon;
        mon.evaluateAtVertices(X, F, M);
        std::cout.precision(17);
        for (int i = 0; i < 3; i++) std::cout << "mon " << M(i, 0) << " " << M(i, 1) << " " << M(i, 2) << " " << M(i, 3) << "\n";
        // the anisotropic form of MEx2: lam1 v v^T + lam2 w w^T
        Eigen::Vector<double, 2> v, w;
        v << 1.0 / sqrt(2.0), 1.0 / sqrt(2.0);
        w << 1.0 / sqrt(2.0), -1.0 / sqrt(2.0);
        Eigen::Matrix<double, 2, 2> m1(3.0 * v * v.transpose()), m2(0.5 * w * w.transpose()), S;
        S = m1 + m2;
        std::cout << "aniso " << S(0, 0) << " " << S(0, 1) << " " << S(1, 0) << " " << S(1, 1) << " det " << S.determinant() << "\n";
    }
    {   // HuangFunctional's virtual pieces (src/HuangFunctional.cpp:35-90): dGdJ against a finite difference of G in J
        HuangFunctional<2> hf(nullptr, 3.5, 0.0, 0.0);
        Eigen::Matrix<double, 2, 2> J, Minv, dJ, dM;
        Eigen::Vector<double, 2> x, dX;
        J << 1.3, 0.2, -0.1, 0.9;
        Minv << 0.7, 0.1, 0.1, 1.4;
        x << 0.3, 0.4;
        double det = J.determinant();
        double G0 = hf.evalG(J, det, Minv, x);
        hf.evaldGdJ(J, det, Minv, x, dJ);
        hf.evaldGdM(J, det, Minv, x, dM);
        hf.evaldGdX(J, det, Minv, x, dX);
        std::cout.precision(17);
        std::cout << "huang G " << G0 << " dGddet " << hf.evaldGddet(J, det, Minv, x) << " dX " << dX(0) << " " << dX(1) << "\n";
        std::cout << "huang dGdJ " << dJ(0, 0) << " " << dJ(0, 1) << " " << dJ(1, 0) << " " << dJ(1, 1) << "\n";
        std::cout << "huang dGdM " << dM(0, 0) << " " << dM(0, 1) << " " << dM(1, 0) << " " << dM(1, 1) << "\n";
        // finite difference of the trace term in J at FIXED det (dGdJ is the partial derivative with det held fixed); out(j,i) = dG/dJ(i,j)
        std::cout << "huang fdJ";
        for (int j = 0; j < 2; j++) for (int i = 0; i < 2; i++) {
            Eigen::Matrix<double, 2, 2> Jp(J); Jp(i, j) += 1e-6;
            Eigen::Matrix<double, 2, 2> Jm(J); Jm(i, j) -= 1e-6;
            std::cout << " " << (hf.evalG(Jp, det, Minv, x) - hf.evalG(Jm, det, Minv, x)) / 2e-6;
        }
        std::cout << "\n";
    }
    {   // MeshIO: the parallel text writer gives the bytes of the reference's ostream loops; binary round trip; readers
        const long R = 50000;
        Eigen::MatrixXd V(R, 2); Eigen::MatrixXi F(R, 3);
        for (long i = 0; i < R; i++) {
            V(i, 0) = (i % 7 == 0) ? 0.0 : sin(0.37 * i) * pow(10.0, (double)(i % 13) - 6); V(i, 1) = -1.0 / (3.0 + i);
            F(i, 0) = (int)i; F(i, 1) = (int)(i * 7 % 1000003); F(i, 2) = -(int)(i % 5);
        }
        {
            std::ofstream out(dir + "/ref_points.txt");
            for (int i = 0; i < V.rows(); i++) { for (int j = 0; j < 1; j++) out << V(i, j) << ", "; out << V(i, 1) << std::endl; }
            std::ofstream out2(dir + "/ref_tris.txt");
            for (int i = 0; i < F.rows(); i++) { for (int j = 0; j < 2; j++) out2 << F(i, j) << ", "; out2 << F(i, 2) << std::endl; }
        }
        std::function<double(long, int)> gv = [&](long i, int j) { return V(i, j); };
        std::function<int(long, int)> gf = [&](long i, int j) { return F(i, j); };
        bool ok = meshio::write_text<double>((dir + "/new_points.txt").c_str(), R, 2, gv) && meshio::write_text<int>((dir + "/new_tris.txt").c_str(), R, 3, gf) &&
                  meshio::write_bin<double>((dir + "/new_points.bin").c_str(), R, 2, gv) && meshio::write_bin<int>((dir + "/new_tris.bin").c_str(), R, 3, gf);
        std::vector<char> a, b, c, d2;
        meshio::slurp((dir + "/ref_points.txt").c_str(), a); meshio::slurp((dir + "/new_points.txt").c_str(), b);
        meshio::slurp((dir + "/ref_tris.txt").c_str(), c); meshio::slurp((dir + "/new_tris.txt").c_str(), d2);
        std::cout << "meshio text " << ((ok && a == b && c == d2) ? "identical" : "DIFFERENT") << "\n";
        std::vector<double> pv; std::vector<int> tv;
        bool rb = meshio::read_doubles((dir + "/new_points.bin").c_str(), pv) && meshio::read_ints((dir + "/new_tris.bin").c_str(), tv);
        bool exact = rb && (long)pv.size() == 2 * R && (long)tv.size() == 3 * R;
        for (long i = 0; exact && i < R; i++) exact = pv[2 * i] == V(i, 0) && pv[2 * i + 1] == V(i, 1) && tv[3 * i + 1] == F(i, 1) && tv[3 * i + 2] == F(i, 2);
        std::cout << "meshio bin " << (exact ? "exact" : "MISMATCH") << "\n";
        // readTriangles from text and from binary (mask: P + 1 entries, reference quirk)
        { std::ofstream m(dir + "/mask.txt"); for (long i = 0; i < R; i++) m << (i % 3) << "\n"; }
        Eigen::MatrixXi F1, F2; Eigen::MatrixXd V1, V2; std::vector<NodeType> m1, m2;
        utils::readTriangles(2, (dir + "/new_tris.txt").c_str(), (dir + "/new_points.txt").c_str(), (dir + "/mask.txt").c_str(), F1, V1, m1);
        utils::readTriangles(2, (dir + "/new_tris.bin").c_str(), (dir + "/new_points.bin").c_str(), (dir + "/mask.txt").c_str(), F2, V2, m2);
        bool rt = F1.rows() == R && V1.rows() == R && (long)m1.size() == R + 1 && F2.rows() == R && m1[4] == (NodeType)1;
        for (long i = 0; rt && i < R; i++) rt = F1(i, 1) == F(i, 1) && F2(i, 2) == F(i, 2) && V2(i, 0) == V(i, 0) && fabs(V1(i, 1) - V(i, 1)) <= 1e-5 * fabs(V(i, 1));
        std::cout << "meshio read " << (rt ? "ok" : "MISMATCH") << "\n";
    }
    {   // json
        std::ifstream in(dir + "/in.json");
        nlohmann::json j;
        in >> j;
        j["Method"] = 2;
        int nx = j["nx"]; double dt = j["dt"]; bool cm = j["CompMesh"]; std::string tt(j["TestType"]);
        std::string s; s = j["TestType"];
        std::cout << "json " << nx << " " << dt << " " << cm << " " << tt << " " << s << " " << j.contains("nz") << " " << j.contains("nx") << "\n";
        std::cout << "dump " << j << "\n";
    }
    return 0;
}
