// Element-block partition of a simplicial mesh across the GPUs of one box -- host-side plan.
//
// The reference has no distributed path (OpenMP over simplices only, src/Mesh.cpp:945-948).  The element-local state
// of the ADMM step (z, u, hessInvs, Ih) needs no communication; only the vertex quantities couple partitions:
// D^T(w(z-u)) in the x-update (src/MeshIntegrator.cpp:137,159) and the assembled gradient of the predictor
// (src/Mesh.cpp:582-624).  Rank r owns the contiguous element block [e0, e1); it keeps a private copy of every vertex
// its elements touch.  Vertices touched by more than one rank are "shared": each rank computes the partial sum over its
// own incident elements, the partials are exchanged pairwise, and every sharer adds them in ascending rank order, so
// all copies of a shared vertex stay bit-identical and the result does not depend on message arrival order.
#pragma once
#include <cstdint>
#include <vector>

namespace mmadmm {

struct PartitionPlan {
    int D = 0, P = 0, E = 0, nranks = 1, rank = 0;
    int e0 = 0, e1 = 0;                 // owned element block
    std::vector<int> verts;             // local vertex -> global vertex (ascending)
    std::vector<int> Floc;              // (e1-e0) x (D+1), local vertex ids
    std::vector<int> valence;           // global valence of every local vertex
    std::vector<uint8_t> owner;         // 1 if this rank is the lowest rank touching the vertex
    // shared vertices (ascending local id) and, per neighbour rank (ascending), the list of shared-vertex slots
    std::vector<int> shared;            // local vertex ids
    std::vector<int> nbrRank;           // neighbour ranks, ascending
    std::vector<int> nbrPtr;            // segment offsets into nbrList (size nbrRank.size()+1)
    std::vector<int> nbrList;           // index into `shared`; both sides hold the segment in ascending GLOBAL vertex order
    // per shared vertex: contributions in ascending rank order.  Entry >= 0: position in the receive buffer
    // (same layout as nbrList); entry == -1: this rank's own partial.
    std::vector<int> contribPtr;        // size shared.size()+1
    std::vector<int> contrib;
};

// Block boundaries: elements split evenly, boundaries rounded to multiples of `align` (whole prox tiles).
void partition_bounds(int E, int nranks, int align, std::vector<int> &bounds);

// Builds the plan of `rank` from the GLOBAL connectivity (every rank holds the global mesh, as every process of the
// reference would).  O(E) host work.  Returns false on bad arguments.
bool build_partition_plan(int D, int P, int E, const int *F, int nranks, int rank, PartitionPlan &plan);

}  // namespace mmadmm
