// Micro-benchmarks that decide the walk kernel's layout on B200 (sm_100a):
//  * LDS.128 wavefront cost: warp-broadcast vs one address per quarter-warp vs per-lane replicated table
//  * fma.rn.f32x2 (FFMA2) vs scalar FFMA issue/throughput, alone and mixed with MUFU/LDS
// Prints cycles per warp-instruction per SM (all 148 SMs busy, WARPS warps per SM).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define ITERS 4096

__device__ __forceinline__ float4 lds128(unsigned a) { float4 v; asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a)); return v; }
__device__ __forceinline__ float2 lds64(unsigned a) { float2 v; asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(a)); return v; }
__device__ __forceinline__ float lds32(unsigned a) { float v; asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a)); return v; }

// mode: address pattern
template <int MODE>
__global__ void k_lds(float * out, long long * cyc)
{
    extern __shared__ __align__(16) unsigned char sm[];
    for(int i = threadIdx.x; i < 65536 / 4; i += blockDim.x) ((float *) sm)[i] = (float) i;
    __syncthreads();
    const unsigned base = (unsigned) __cvta_generic_to_shared(sm);
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned a = base + warp * 2048;
    unsigned stride = 16;
    if(MODE == 0) a += 0;                       // LDS.128 all lanes same address
    if(MODE == 1) a += (lane >> 3) * 16;        // LDS.128 one address per quarter warp (consecutive 16B)
    if(MODE == 2) a += (lane >> 3) * 528;       // LDS.128 one address per quarter warp (far apart, distinct bank groups)
    if(MODE == 3) a += (lane & 3) * 16;         // LDS.128 4 addresses interleaved over lanes
    if(MODE == 4) a += lane * 16;               // LDS.128 all distinct, conflict-free (4 wavefronts)
    if(MODE == 5) a += 0;                       // LDS.64 same address
    if(MODE == 6) a += (lane & 15) * 8;         // LDS.64 replicated-table pattern: lane picks copy (lane&15), random row
    if(MODE == 7) a += 0;                       // LDS.64 random row of a single copy (conflicts)
    if(MODE == 8) a += (lane >> 3) * 4;         // LDS.32 4 addresses distinct banks
    if(MODE == 9) a += (lane >> 3) * 16 ;       // LDS.128, k advancing in lockstep (lists at +q*16): like mode 1
    if(MODE == 10) a += (lane & 7) * 16;        // LDS.128 replicated POT table: copy lane&7, random row
    float acc = 0.f;
    unsigned r = lane * 2654435761u + warp * 40503u + 12345u;
    long long t0 = clock64();
    #pragma unroll 8
    for(int i = 0; i < ITERS; i++) {
        unsigned aa = a + ((i & 63) * stride);
        if(MODE == 6) { r = r * 1664525u + 1013904223u; aa = a - warp * 2048 + ((r >> 23) & 511) * 128; }
        if(MODE == 7) { r = r * 1664525u + 1013904223u; aa = a - warp * 2048 + ((r >> 23) & 511) * 8; }
        if(MODE == 10) { r = r * 1664525u + 1013904223u; aa = a - warp * 2048 + ((r >> 23) & 511) * 128; }
        if(MODE <= 4 || MODE == 9 || MODE == 10) { float4 v = lds128(aa); acc += v.x + v.w; }
        else if(MODE <= 7) { float2 v = lds64(aa); acc += v.x + v.y; }
        else { acc += lds32(aa); }
    }
    long long t1 = clock64();
    if(acc == 1234.5f) out[0] = acc;
    if(threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

// FMA throughput: scalar FFMA vs packed f32x2
template <int MODE>
__global__ void k_fma(float * out, long long * cyc)
{
    float x0 = threadIdx.x * 1e-3f, x1 = x0 + 1.f, x2 = x0 + 2.f, x3 = x0 + 3.f, x4 = x0 + 4.f, x5 = x0 + 5.f, x6 = x0 + 6.f, x7 = x0 + 7.f;
    const float a = 1.0001f, b = 1e-4f;
    long long t0 = clock64();
    if(MODE == 0) {
        #pragma unroll 4
        for(int i = 0; i < ITERS; i++) {
            x0 = fmaf(x0, a, b); x1 = fmaf(x1, a, b); x2 = fmaf(x2, a, b); x3 = fmaf(x3, a, b);
            x4 = fmaf(x4, a, b); x5 = fmaf(x5, a, b); x6 = fmaf(x6, a, b); x7 = fmaf(x7, a, b);
        }
    } else {
        unsigned long long p0, p1, p2, p3, pa, pb;
        asm("mov.b64 %0, {%1, %2};" : "=l"(p0) : "f"(x0), "f"(x1));
        asm("mov.b64 %0, {%1, %2};" : "=l"(p1) : "f"(x2), "f"(x3));
        asm("mov.b64 %0, {%1, %2};" : "=l"(p2) : "f"(x4), "f"(x5));
        asm("mov.b64 %0, {%1, %2};" : "=l"(p3) : "f"(x6), "f"(x7));
        asm("mov.b64 %0, {%1, %2};" : "=l"(pa) : "f"(a), "f"(a));
        asm("mov.b64 %0, {%1, %2};" : "=l"(pb) : "f"(b), "f"(b));
        #pragma unroll 4
        for(int i = 0; i < ITERS; i++) {
            asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p0) : "l"(pa), "l"(pb));
            asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p1) : "l"(pa), "l"(pb));
            asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p2) : "l"(pa), "l"(pb));
            asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p3) : "l"(pa), "l"(pb));
            if(MODE == 2) {   // interleave ALU work that would compete for issue slots
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(*(unsigned *) &x0) : "r"(i), "r"(0x55));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(*(unsigned *) &x1) : "r"(i), "r"(0x33));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(*(unsigned *) &x2) : "r"(i), "r"(0x0f));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(*(unsigned *) &x3) : "r"(i), "r"(0xff));
            }
        }
        asm("mov.b64 {%0, %1}, %2;" : "=f"(x0), "=f"(x1) : "l"(p0));
        float y0, y1;
        asm("mov.b64 {%0, %1}, %2;" : "=f"(y0), "=f"(y1) : "l"(p1)); x2 += y0 + y1;
        asm("mov.b64 {%0, %1}, %2;" : "=f"(y0), "=f"(y1) : "l"(p2)); x4 += y0 + y1;
        asm("mov.b64 {%0, %1}, %2;" : "=f"(y0), "=f"(y1) : "l"(p3)); x6 += y0 + y1;
    }
    long long t1 = clock64();
    float s = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
    if(s == 1234.5f) out[0] = s;
    if(threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <class F> void run(const char * name, F kern, int threads, size_t smem, double instr_per_iter_per_warp)
{
    float * out; long long * cyc;
    cudaMalloc(&out, 1024); cudaMalloc(&cyc, 148 * 8);
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
    kern<<<148, threads, smem>>>(out, cyc);
    kern<<<148, threads, smem>>>(out, cyc);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[148];
    cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0; for(int i = 0; i < 148; i++) avg += h[i]; avg /= 148;
    const double winstr = (double) ITERS * instr_per_iter_per_warp * (threads / 32);
    printf("%-58s %s cycles %.0f  -> %.3f clk per warp-instr per SM\n", name, cudaGetErrorString(e), avg, avg / winstr);
    cudaFree(out); cudaFree(cyc);
}

int main()
{
    const int T = 512; const size_t S = 65536 + 2048 * 16;
    run("LDS.128 warp broadcast", k_lds<0>, T, S, 1);
    run("LDS.128 one addr per quarter-warp, adjacent 16B", k_lds<1>, T, S, 1);
    run("LDS.128 one addr per quarter-warp, distinct bank groups", k_lds<2>, T, S, 1);
    run("LDS.128 4 addrs interleaved (lane&3)", k_lds<3>, T, S, 1);
    run("LDS.128 32 distinct conflict-free", k_lds<4>, T, S, 1);
    run("LDS.64 warp broadcast", k_lds<5>, T, S, 1);
    run("LDS.64 replicated table (copy lane&15), random rows", k_lds<6>, T, S, 1);
    run("LDS.64 single table random rows", k_lds<7>, T, S, 1);
    run("LDS.32 one addr per quarter-warp", k_lds<8>, T, S, 1);
    run("LDS.128 replicated POT table (copy lane&7), random rows", k_lds<10>, T, S, 1);
    run("FFMA scalar x8", k_fma<0>, T, 0, 8);
    run("FFMA2 packed x4 (8 fma)", k_fma<1>, T, 0, 4);
    run("FFMA2 x4 + LOP3 x4", k_fma<2>, T, 0, 8);
    return 0;
}
