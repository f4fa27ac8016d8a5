// microbench.cu — the roofs the render kernels are compared against, measured on the
// box the bench runs on:   nvcc -gencode arch=compute_100a,code=sm_100a -O3 tools/microbench.cu -o /tmp/mb && /tmp/mb
//   l2_read_gbs      grid-stride float4 reads over a 32 MB buffer (L2 resident after the first pass)
//   hbm_read_gbs     the same over 4 GB
//   issue_ginst      independent integer ops, warp instructions per second (roof: SMs x 4 schedulers x clock)
//   fp64 / fp32      DFMA, DADD, DMUL, FFMA warp instructions per second, 8 independent chains per thread
//   l2_chase_ns      dependent-load latency over 32 MB (one thread), dependent 64-B node fetch as the BVH walk does
// One JSON line on stdout.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x)                                                                                  \
    do {                                                                                       \
        cudaError_t e_ = (x);                                                                  \
        if (e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } \
    } while (0)

__global__ void readKernel(const float4* __restrict__ p, size_t n, int reps, float* out) {
    float acc = 0.f;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (int r = 0; r < reps; r++)
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
            const float4 v = __ldg(p + i);
            acc += v.x + v.y + v.z + v.w;
        }
    if (acc == 123.456f) *out = acc;
}

template <int OP>
__global__ void aluKernel(int iters, double* out, double seed) {
    double a[8];
    float f[8];
    unsigned u[8];
    for (int k = 0; k < 8; k++) { a[k] = seed + k; f[k] = (float)seed + k; u[k] = (unsigned)k + threadIdx.x; }
    const double m = 1.0000001, c = 1e-9;
    const float mf = 1.0000001f, cf = 1e-9f;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < 8; k++) {
            if (OP == 0) a[k] = fma(a[k], m, c);
            if (OP == 1) a[k] = __dadd_rn(a[k], c);
            if (OP == 2) a[k] = __dmul_rn(a[k], m);
            if (OP == 3) f[k] = __fmaf_rn(f[k], mf, cf);
            if (OP == 4) u[k] = (u[k] ^ (unsigned)i) + 0x9e3779b9u;     // LOP3 + IADD3: two instructions
        }
    }
    double s = 0;
    for (int k = 0; k < 8; k++) s += a[k] + f[k] + u[k];
    if (s == 123.456) *out = s;
}

__global__ void chaseKernel(const int* __restrict__ next, int start, int steps, int* out) {
    int cur = start;
    for (int i = 0; i < steps; i++) cur = __ldg(next + (size_t)cur * 16);     // one 64-B record per hop
    *out = cur;
}

static float timeIt(cudaEvent_t e0, cudaEvent_t e1) {
    float ms;
    CK(cudaEventSynchronize(e1));
    CK(cudaEventElapsedTime(&ms, e0, e1));
    return ms;
}

int main() {
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    const int sms = prop.multiProcessorCount;
    int clockKHz = 0;
    CK(cudaDeviceGetAttribute(&clockKHz, cudaDevAttrClockRate, 0));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float* dOut; CK(cudaMalloc(&dOut, 64));

    auto readBw = [&](size_t bytes, int reps) {
        float4* p; CK(cudaMalloc(&p, bytes)); CK(cudaMemset(p, 0, bytes));
        const size_t n = bytes / 16;
        double best = 0;
        for (int t = 0; t < 5; t++) {
            CK(cudaEventRecord(e0));
            readKernel<<<sms * 8, 256>>>(p, n, reps, dOut);
            CK(cudaEventRecord(e1));
            const float ms = timeIt(e0, e1);
            const double gbs = (double)bytes * reps / (ms * 1e-3) / 1e9;
            if (t > 0 && gbs > best) best = gbs;
        }
        CK(cudaFree(p));
        return best;
    };
    const double l2 = readBw(32u << 20, 200);
    const double l2b = readBw(96u << 20, 60);
    const double hbm = readBw((size_t)4 << 30, 2);

    auto alu = [&](int op) {
        const int iters = 4096, blocks = sms * 8, threads = 256;
        double best = 0;
        for (int t = 0; t < 4; t++) {
            CK(cudaEventRecord(e0));
            switch (op) {
                case 0: aluKernel<0><<<blocks, threads>>>(iters, (double*)dOut, 1.0); break;
                case 1: aluKernel<1><<<blocks, threads>>>(iters, (double*)dOut, 1.0); break;
                case 2: aluKernel<2><<<blocks, threads>>>(iters, (double*)dOut, 1.0); break;
                case 3: aluKernel<3><<<blocks, threads>>>(iters, (double*)dOut, 1.0); break;
                default: aluKernel<4><<<blocks, threads>>>(iters, (double*)dOut, 1.0); break;
            }
            CK(cudaEventRecord(e1));
            const float ms = timeIt(e0, e1);
            const double warpInst = (double)blocks * (threads / 32) * iters * 8 * (op == 4 ? 2 : 1);
            const double g = warpInst / (ms * 1e-3) / 1e9;
            if (t > 0 && g > best) best = g;
        }
        return best;       // G warp-instructions / s
    };
    const double dfma = alu(0), dadd = alu(1), dmul = alu(2), ffma = alu(3), iop = alu(4);

    // dependent 64-B record chase over 32 MB: a random cyclic permutation
    const int nRec = (32 << 20) / 64;
    std::vector<int> perm(nRec), nxt((size_t)nRec * 16, 0);
    for (int i = 0; i < nRec; i++) perm[i] = i;
    unsigned long long s = 88172645463325252ULL;
    for (int i = nRec - 1; i > 0; i--) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; const int j = (int)(s % (unsigned)(i + 1)); std::swap(perm[i], perm[j]); }
    for (int i = 0; i < nRec; i++) nxt[(size_t)perm[i] * 16] = perm[(i + 1) % nRec];
    int* dNext; CK(cudaMalloc(&dNext, nxt.size() * 4));
    CK(cudaMemcpy(dNext, nxt.data(), nxt.size() * 4, cudaMemcpyHostToDevice));
    const int steps = 200000;
    chaseKernel<<<1, 1>>>(dNext, 0, steps, (int*)dOut);       // warm L2
    CK(cudaEventRecord(e0));
    chaseKernel<<<1, 1>>>(dNext, perm[7], steps, (int*)dOut);
    CK(cudaEventRecord(e1));
    const double chaseNs = timeIt(e0, e1) * 1e6 / steps;

    const double issueRoof = (double)sms * 4 * clockKHz * 1e3 / 1e9;
    printf("{\"gpu\": \"%s\", \"sms\": %d, \"clock_mhz\": %.0f, \"l2_read_gbs_32mb\": %.1f, \"l2_read_gbs_96mb\": %.1f, \"hbm_read_gbs_4gb\": %.1f, "
           "\"issue_roof_ginst\": %.1f, \"int_ginst\": %.1f, \"dfma_ginst\": %.2f, \"dadd_ginst\": %.2f, \"dmul_ginst\": %.2f, \"ffma_ginst\": %.1f, "
           "\"fp64_gflops_fma\": %.0f, \"fp64_gflops_nofma\": %.0f, \"fp32_gflops_fma\": %.0f, \"l2_chase_ns_64B\": %.1f}\n",
           prop.name, sms, clockKHz / 1e3, l2, l2b, hbm, issueRoof, iop, dfma, dadd, dmul, ffma, dfma * 32 * 2, dadd * 32, ffma * 32 * 2, chaseNs);
    return 0;
}
