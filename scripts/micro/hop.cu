// Microbenchmark: what one producer->consumer hop through L2 costs on this GPU (st.relaxed.gpu -> ld.relaxed.gpu polling),
// and the latency of dependent ld.relaxed.gpu / ld.global.cg / plain loads that hit L2.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o hop hop.cu
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ unsigned long long ld_rel(const unsigned long long *p) { unsigned long long v; asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ void st_rel(unsigned long long *p, unsigned long long v) { asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory"); }
__device__ __forceinline__ unsigned long long gt() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
// ping-pong: CTA 0 and CTA `other` bounce a counter `iters` times through two flags (one thread each)
__global__ void k_pingpong(unsigned long long *flags, int other, int iters, unsigned long long *ns) {
    if (threadIdx.x != 0) return;
    if (blockIdx.x == 0) {
        unsigned long long t0 = gt();
        for (int i = 1; i <= iters; i++) { st_rel(flags, i); while (ld_rel(flags + 32) != (unsigned long long)i) {} }
        *ns = gt() - t0;
    } else if (blockIdx.x == other) {
        for (int i = 1; i <= iters; i++) { while (ld_rel(flags) != (unsigned long long)i) {} st_rel(flags + 32, i); }
    }
}
// chain: a relay of `len` CTAs, each waits for its predecessor's value and publishes value+1 (a wavefront of width 1)
__global__ void k_relay(unsigned long long *cells, int len, unsigned long long *ns) {
    if (threadIdx.x != 0) return;
    const int b = blockIdx.x;
    unsigned long long t0 = gt();
    if (b == 0) { st_rel(cells, 1); }
    else { unsigned long long v; while ((v = ld_rel(cells + 32 * (b - 1))) == 0) {} st_rel(cells + 32 * b, v + 1); }
    if (b == len - 1) *ns = gt() - t0;
}
template <int MODE> __global__ void k_chase(const unsigned long long *p, int iters, unsigned long long *out, unsigned long long *ns) {
    unsigned long long i = 0, t0 = gt();
    for (int k = 0; k < iters; k++) {
        if (MODE == 0) i = ld_rel(p + i);
        else if (MODE == 1) i = __ldcg(p + i);
        else i = p[i];
    }
    *ns = gt() - t0; *out = i;
}
int main() {
    unsigned long long *flags, *ns, *buf, h;
    cudaMalloc(&flags, 1 << 20); cudaMalloc(&ns, 64); cudaMemset(flags, 0, 1 << 20);
    int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    for (int other : {1, 2, 37, 74, 75, 110, 147}) {
        if (other >= sms) continue;
        cudaMemset(flags, 0, 1 << 20);
        k_pingpong<<<sms, 32>>>(flags, other, 2000, ns);
        cudaMemcpy(&h, ns, 8, cudaMemcpyDeviceToHost);
        printf("pingpong CTA0 <-> CTA%-3d: %.3f us per round trip = %.3f us per hop\n", other, h / 2000.0 / 1e3, h / 4000.0 / 1e3);
    }
    for (int len : {16, 64, 148}) {
        cudaMemset(flags, 0, 1 << 20);
        k_relay<<<len, 32>>>(flags, len, ns);
        cudaMemcpy(&h, ns, 8, cudaMemcpyDeviceToHost);
        printf("relay over %3d CTAs: %.3f us per hop (includes launch skew)\n", len, h / (double)(len - 1) / 1e3);
    }
    const int N = 1 << 17;   // 1 MiB of indices: L2 resident
    unsigned long long *hb = new unsigned long long[N];
    for (int i = 0; i < N; i++) hb[i] = (i * 40503ull + 12345) % N;
    cudaMalloc(&buf, N * 8); cudaMemcpy(buf, hb, N * 8, cudaMemcpyHostToDevice);
    unsigned long long *o; cudaMalloc(&o, 8);
    k_chase<0><<<1, 1>>>(buf, 20000, o, ns); cudaMemcpy(&h, ns, 8, cudaMemcpyDeviceToHost); printf("dependent ld.relaxed.gpu: %.1f ns per load\n", h / 20000.0);
    k_chase<1><<<1, 1>>>(buf, 20000, o, ns); cudaMemcpy(&h, ns, 8, cudaMemcpyDeviceToHost); printf("dependent ld.global.cg : %.1f ns per load\n", h / 20000.0);
    k_chase<2><<<1, 1>>>(buf, 20000, o, ns); cudaMemcpy(&h, ns, 8, cudaMemcpyDeviceToHost); printf("dependent plain ld     : %.1f ns per load\n", h / 20000.0);
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
