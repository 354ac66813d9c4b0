// Host-side element-block partition plan (see partition.h).  No device code: usable (and tested) without a GPU.
#include "partition.h"

#include <algorithm>

namespace mmadmm {

void partition_bounds(int E, int nranks, int align, std::vector<int> &bounds) {
    bounds.assign(nranks + 1, 0);
    for (int r = 1; r < nranks; r++) {
        long long b = (long long)E * r / nranks;
        b = (b + align / 2) / align * align;
        if (b > E) b = E;
        if (b < bounds[r - 1]) b = bounds[r - 1];
        bounds[r] = (int)b;
    }
    bounds[nranks] = E;
}

bool build_partition_plan(int D, int P, int E, const int *F, int nranks, int rank, PartitionPlan &plan) {
    if ((D != 2 && D != 3) || P <= 0 || E <= 0 || !F || nranks < 1 || nranks > 64 || rank < 0 || rank >= nranks) return false;
    const int V = D + 1;
    plan = PartitionPlan();
    plan.D = D; plan.P = P; plan.E = E; plan.nranks = nranks; plan.rank = rank;
    std::vector<int> bounds;
    partition_bounds(E, nranks, 128, bounds);
    plan.e0 = bounds[rank];
    plan.e1 = bounds[rank + 1];

    // which ranks touch each vertex, and the global valence
    std::vector<uint64_t> touch((size_t)P, 0);
    std::vector<int> val((size_t)P, 0);
    for (int r = 0; r < nranks; r++) {
        const uint64_t bit = 1ull << r;
        for (size_t i = (size_t)bounds[r] * V; i < (size_t)bounds[r + 1] * V; i++) {
            const int v = F[i];
            if (v < 0 || v >= P) return false;
            touch[v] |= bit;
            val[v]++;
        }
    }
    const uint64_t me = 1ull << rank;
    std::vector<int> g2l((size_t)P, -1);
    for (int v = 0; v < P; v++)
        if (touch[v] & me) {
            g2l[v] = (int)plan.verts.size();
            plan.verts.push_back(v);
        }
    const int Pl = (int)plan.verts.size();
    plan.valence.resize(Pl);
    plan.owner.resize(Pl);
    std::vector<int> sharedSlot((size_t)Pl, -1);
    for (int l = 0; l < Pl; l++) {
        const int v = plan.verts[l];
        plan.valence[l] = val[v];
        plan.owner[l] = ((touch[v] & (me - 1)) == 0) ? 1 : 0;   // no lower rank touches it
        if (touch[v] != me) {
            sharedSlot[l] = (int)plan.shared.size();
            plan.shared.push_back(l);
        }
    }
    plan.Floc.resize((size_t)(plan.e1 - plan.e0) * V);
    for (size_t i = 0; i < plan.Floc.size(); i++) plan.Floc[i] = g2l[F[(size_t)plan.e0 * V + i]];

    // neighbour segments (ascending neighbour rank; ascending global vertex inside a segment)
    const int nSh = (int)plan.shared.size();
    std::vector<std::vector<std::pair<int, int>>> perVertex((size_t)nSh);   // (rank, receive position)
    plan.nbrPtr.push_back(0);
    for (int q = 0; q < nranks; q++) {
        if (q == rank) {
            for (int i = 0; i < nSh; i++) perVertex[i].push_back({q, -1});
            continue;
        }
        const uint64_t qb = 1ull << q;
        const size_t before = plan.nbrList.size();
        for (int i = 0; i < nSh; i++) {
            const int v = plan.verts[plan.shared[i]];
            if (touch[v] & qb) {
                perVertex[i].push_back({q, (int)plan.nbrList.size()});
                plan.nbrList.push_back(i);
            }
        }
        if (plan.nbrList.size() > before) {
            plan.nbrRank.push_back(q);
            plan.nbrPtr.push_back((int)plan.nbrList.size());
        }
    }
    plan.contribPtr.push_back(0);
    for (int i = 0; i < nSh; i++) {
        for (auto &pr : perVertex[i]) plan.contrib.push_back(pr.second);
        plan.contribPtr.push_back((int)plan.contrib.size());
    }
    return true;
}

}  // namespace mmadmm

// ---- C ABI (include/mmadmm.h): plan inspection, usable without a device ----
#include "../../include/mmadmm.h"

struct mmadmm_plan { mmadmm::PartitionPlan p; };

extern "C" {

int mmadmm_plan_create(int D, int P, int E, const int *F, int nranks, int rank, mmadmm_plan **out) {
    if (!out) return MMADMM_EINVAL;
    *out = nullptr;
    mmadmm_plan *pl = new mmadmm_plan();
    if (!mmadmm::build_partition_plan(D, P, E, F, nranks, rank, pl->p)) { delete pl; return MMADMM_EINVAL; }
    *out = pl;
    return MMADMM_OK;
}

int mmadmm_plan_sizes(const mmadmm_plan *pl, int *sizes6) {
    if (!pl || !sizes6) return MMADMM_EINVAL;
    sizes6[0] = pl->p.e0; sizes6[1] = pl->p.e1; sizes6[2] = (int)pl->p.verts.size();
    sizes6[3] = (int)pl->p.shared.size(); sizes6[4] = (int)pl->p.nbrRank.size(); sizes6[5] = (int)pl->p.nbrList.size();
    return MMADMM_OK;
}

int mmadmm_plan_get(const mmadmm_plan *pl, int which, int *out) {
    if (!pl || !out) return MMADMM_EINVAL;
    const std::vector<int> *v = nullptr;
    switch (which) {
        case MMADMM_PLAN_VERTS: v = &pl->p.verts; break;
        case MMADMM_PLAN_FLOC: v = &pl->p.Floc; break;
        case MMADMM_PLAN_VALENCE: v = &pl->p.valence; break;
        case MMADMM_PLAN_SHARED: v = &pl->p.shared; break;
        case MMADMM_PLAN_NBR_RANK: v = &pl->p.nbrRank; break;
        case MMADMM_PLAN_NBR_PTR: v = &pl->p.nbrPtr; break;
        case MMADMM_PLAN_NBR_LIST: v = &pl->p.nbrList; break;
        case MMADMM_PLAN_CONTRIB_PTR: v = &pl->p.contribPtr; break;
        case MMADMM_PLAN_CONTRIB: v = &pl->p.contrib; break;
        case MMADMM_PLAN_OWNER:
            for (size_t i = 0; i < pl->p.owner.size(); i++) out[i] = pl->p.owner[i];
            return MMADMM_OK;
        default: return MMADMM_EINVAL;
    }
    std::copy(v->begin(), v->end(), out);
    return MMADMM_OK;
}

void mmadmm_plan_destroy(mmadmm_plan *pl) { delete pl; }

}  // extern "C"
