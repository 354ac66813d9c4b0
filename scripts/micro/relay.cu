// Microbenchmark: a relay of warps shaped like one column of the dataflow triangular solve (k_tri_sf_block).  Warp w waits until
// the 32 two-component results of warp w-1 (scattered `stride` doubles apart, pre-filled with a sentinel) are visible, runs a
// dependent FP64 chain of `nfma` FMAs and `ndiv` divisions per lane, and publishes its own 32 results.  Prints us per hop.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o relay relay.cu
#include <cstdio>
#include <cuda_runtime.h>
constexpr unsigned long long SENT = 0xFFF8DEADBEEF0001ull;
__device__ __forceinline__ unsigned long long gt() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
__device__ __forceinline__ bool ld2(const double *p, double &a, double &b) {
    asm volatile("ld.relaxed.gpu.global.v2.f64 {%0, %1}, [%2];" : "=d"(a), "=d"(b) : "l"(p) : "memory");
    return (unsigned long long)__double_as_longlong(a) != SENT && (unsigned long long)__double_as_longlong(b) != SENT;
}
__device__ __forceinline__ void st2(double *p, double a, double b) { asm volatile("st.relaxed.gpu.global.v2.f64 [%0], {%1, %2};" ::"l"(p), "d"(a), "d"(b) : "memory"); }
// mode bit 0: proxy phase (lane 0 polls first, then everybody); bit 1: two neighbours polled one after the other (else one)
// mode bit 2: results of consecutive warps ADJACENT in memory (lane stripes of `stride` doubles, warp w at offset 2 w inside its lane's
// stripe): a 32-byte sector then holds the results of two consecutive levels, as in the solve (vertices v and v + 1)
__global__ void k_relay(double *cells, size_t stride, int nwarps, int warpsPerCta, int nfma, int ndiv, int mode, unsigned long long *ns) {
    const int lane = threadIdx.x & 31, w = blockIdx.x * warpsPerCta + (threadIdx.x >> 5);
    if (w >= nwarps) return;
    if (mode & 4) {
        double *mine = cells + (size_t)lane * stride + 2 * (size_t)w;
        unsigned long long t0 = gt();
        double a = 1.0, b = 1.0;
        if (w > 0) { while (!ld2(mine - 2, a, b)) {} __syncwarp(); }
        double s = a;
        for (int k = 0; k < nfma; k++) s = fma(-1e-9, b, s);
        for (int k = 0; k < ndiv; k++) s = s / 1.0000001;
        st2(mine, s, s + 1.0);
        if (w == nwarps - 1 && lane == 0) *ns = gt() - t0;
        return;
    }
    double *mine = cells + ((size_t)w * 32 + lane) * stride;
    unsigned long long t0 = gt();
    double a = 1.0, b = 1.0;
    if (w > 0) {
        const double *prev = cells + ((size_t)(w - 1) * 32 + lane) * stride, *prev2 = cells + ((size_t)(w - 1) * 32 + ((lane + 1) & 31)) * stride;
        if (mode & 1) {
            if (lane == 0) { double x, y; while (!ld2(prev, x, y)) {} }
            __syncwarp();
        }
        while (!ld2(prev, a, b)) {}
        if (mode & 2) { double c, d; while (!ld2(prev2, c, d)) {} a += 1e-30 * c; b += 1e-30 * d; }
        __syncwarp();
    }
    double s = a;
    for (int k = 0; k < nfma; k++) s = fma(-1e-9, b, s);
    for (int k = 0; k < ndiv; k++) s = s / 1.0000001;
    st2(mine, s, s + 1.0);
    if (w == nwarps - 1 && lane == 0) *ns = gt() - t0;
}
int main() {
    const int nwarps = 1184;
    const size_t stride = 4096;   // doubles between the results of neighbouring lanes (a lattice row apart: 2049 vertices x 2)
    const size_t n = (size_t)nwarps * 32 * stride;
    double *cells; unsigned long long *ns, h;
    cudaMalloc(&cells, n * 8); cudaMalloc(&ns, 8);
    unsigned long long *hs = new unsigned long long[n];
    for (size_t i = 0; i < n; i++) hs[i] = SENT;
    for (int wpc : {1, 8}) for (int mode : {0, 1, 2, 3}) for (int work : {0, 1}) {
        cudaMemcpy(cells, hs, n * 8, cudaMemcpyHostToDevice);
        const int nf = work ? 26 : 0, nd = work ? 2 : 0;
        k_relay<<<(nwarps + wpc - 1) / wpc, 32 * wpc>>>(cells, stride, nwarps, wpc, nf, nd, mode, ns);
        cudaMemcpy(&h, ns, 8, cudaMemcpyDeviceToHost);
        printf("warps/CTA %d  proxy %d  two neighbours %d  chain %2d fma + %d div : %.3f us per hop\n", wpc, mode & 1, (mode >> 1) & 1, nf, nd, h / 1e3 / (nwarps - 1));
    }
    for (int wpc : {1, 8}) for (int work : {0, 1}) {
        cudaMemcpy(cells, hs, n * 8, cudaMemcpyHostToDevice);
        const int nf = work ? 26 : 0, nd = work ? 2 : 0;
        k_relay<<<(nwarps + wpc - 1) / wpc, 32 * wpc>>>(cells, stride, nwarps, wpc, nf, nd, 4, ns);
        cudaMemcpy(&h, ns, 8, cudaMemcpyDeviceToHost);
        printf("warps/CTA %d  results of consecutive levels share sectors  chain %2d fma + %d div : %.3f us per hop\n", wpc, nf, nd, h / 1e3 / (nwarps - 1));
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
