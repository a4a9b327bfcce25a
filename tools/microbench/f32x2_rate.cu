#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ float2 add2(float2 a, float2 b){ float2 r; asm("{.reg .b64 ra, rb, rc; mov.b64 ra, {%2,%3}; mov.b64 rb, {%4,%5}; add.rn.f32x2 rc, ra, rb; mov.b64 {%0,%1}, rc;}" : "=f"(r.x), "=f"(r.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y)); return r; }
__device__ __forceinline__ float2 mul2(float2 a, float2 b){ float2 r; asm("{.reg .b64 ra, rb, rc; mov.b64 ra, {%2,%3}; mov.b64 rb, {%4,%5}; mul.rn.f32x2 rc, ra, rb; mov.b64 {%0,%1}, rc;}" : "=f"(r.x), "=f"(r.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y)); return r; }
__device__ __forceinline__ float2 sub2(float2 a, float2 b){ float2 r; asm("{.reg .b64 ra, rb, rc; mov.b64 ra, {%2,%3}; mov.b64 rb, {%4,%5}; sub.rn.f32x2 rc, ra, rb; mov.b64 {%0,%1}, rc;}" : "=f"(r.x), "=f"(r.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y)); return r; }
template<int MODE> __global__ void k(float* out, int iters, float s)
{
    float2 a[4], b[4];
    for (int i = 0; i < 4; i++) { a[i] = make_float2(threadIdx.x * 0.001f + i, threadIdx.x * 0.002f + i); b[i] = make_float2(s, s * 0.5f); }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 4; i++) {
            if (MODE == 0) { a[i].x = __fadd_rn(__fmul_rn(a[i].x, b[i].x), b[i].y); a[i].y = __fadd_rn(__fmul_rn(a[i].y, b[i].x), b[i].y); }
            else { a[i] = add2(mul2(a[i], b[i]), b[i]); }
        }
    }
    float r = 0; for (int i = 0; i < 4; i++) r += a[i].x + a[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
int main(){
    float* d; cudaMalloc(&d, 148*8*256*4);
    cudaEvent_t e0,e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int mode=0; mode<2; mode++) for (int rep=0; rep<3; rep++){
        cudaEventRecord(e0);
        if (mode==0) k<0><<<148*8,256>>>(d, 100000, 1.0001f); else k<1><<<148*8,256>>>(d, 100000, 1.0001f);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms,e0,e1);
        double flops = 148.0*8*256*100000*4*2*2; // per iter per i: 2 lanes x (mul+add)
        printf("mode %d: %.3f ms, %.2f T scalar-flop/s\n", mode, ms, flops/ms/1e9);
    }
    return 0;
}
