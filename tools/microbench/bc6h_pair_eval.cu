// microbench: BC6H nearest-entry error, scalar (current) vs packed-f32x2 over the two perturbation signs (planned)
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
struct f2 { float x, y; };
__device__ __forceinline__ unsigned long long pk(float a, float b){ unsigned long long r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void upk(unsigned long long v, float& a, float& b){ asm("mov.b64 {%0,%1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ unsigned long long sub2(unsigned long long a, unsigned long long b){ unsigned long long r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ unsigned long long sq2(unsigned long long a, unsigned long long z){ unsigned long long r; asm("fma.rn.f32x2 %0, %1, %1, %2;" : "=l"(r) : "l"(a), "l"(z)); return r; }
__device__ __forceinline__ unsigned long long add2(unsigned long long a, unsigned long long b){ unsigned long long r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }

// current: one candidate palette, 8 entries, 16 texels
__global__ void __launch_bounds__(128) k_scalar(const float* __restrict__ tex, const float* __restrict__ pal, float* out, int iters)
{
    __shared__ float st[48 * 128];
    for (int i = 0; i < 48; i++) st[i * 128 + threadIdx.x] = tex[(blockIdx.x * 128 + threadIdx.x) * 48 + i];
    float p[8][3];
    for (int j = 0; j < 8; j++) for (int c = 0; c < 3; c++) p[j][c] = pal[j * 3 + c] + threadIdx.x;
    float acc = 0;
    for (int it = 0; it < iters; it++) {
        float tot = 0;
#pragma unroll 1
        for (int i = 0; i < 16; i++) {
            const float r = st[(3 * i) * 128 + threadIdx.x], g = st[(3 * i + 1) * 128 + threadIdx.x], b = st[(3 * i + 2) * 128 + threadIdx.x];
            float dr = r - p[0][0], dg = g - p[0][1], db = b - p[0][2];
            float best = dr * dr + dg * dg + db * db;
            bool open = true;
#pragma unroll
            for (int j = 1; j < 8; j++) {
                open = open && best > 0.0f;
                dr = r - p[j][0]; dg = g - p[j][1]; db = b - p[j][2];
                const float e = dr * dr + dg * dg + db * db;
                if (open && e > best) open = false;
                if (open && e < best) best = e;
            }
            tot += best;
        }
        acc += tot;
        p[0][0] += 1.0f; // keep the loop from being hoisted
    }
    out[blockIdx.x * 128 + threadIdx.x] = acc;
}
// planned: two candidate palettes (x = minus, y = plus) in f32x2 lanes
__global__ void __launch_bounds__(128) k_packed(const float* __restrict__ tex, const float* __restrict__ pal, float* out, int iters, unsigned long long Z)
{
    __shared__ float st[48 * 128];
    for (int i = 0; i < 48; i++) st[i * 128 + threadIdx.x] = tex[(blockIdx.x * 128 + threadIdx.x) * 48 + i];
    unsigned long long p[8][3];
    for (int j = 0; j < 8; j++) for (int c = 0; c < 3; c++) p[j][c] = pk(pal[j * 3 + c] + threadIdx.x, pal[j * 3 + c] + threadIdx.x + (c == 0 ? 3.0f : 0.0f));
    float acc = 0;
    for (int it = 0; it < iters; it++) {
        float tot0 = 0, tot1 = 0;
#pragma unroll 1
        for (int i = 0; i < 16; i++) {
            const float r = st[(3 * i) * 128 + threadIdx.x], g = st[(3 * i + 1) * 128 + threadIdx.x], b = st[(3 * i + 2) * 128 + threadIdx.x];
            const unsigned long long rr = pk(r, r), gg = pk(g, g), bb = pk(b, b);
            unsigned long long dr = sub2(rr, p[0][0]), dg = sub2(gg, p[0][1]), db = sub2(bb, p[0][2]);
            unsigned long long e2 = add2(add2(sq2(dr, Z), sq2(dg, Z)), sq2(db, Z));
            float b0, b1;
            upk(e2, b0, b1);
            bool o0 = true, o1 = true;
#pragma unroll
            for (int j = 1; j < 8; j++) {
                dr = sub2(rr, p[j][0]); dg = sub2(gg, p[j][1]); db = sub2(bb, p[j][2]);
                e2 = add2(add2(sq2(dr, Z), sq2(dg, Z)), sq2(db, Z));
                float e0, e1;
                upk(e2, e0, e1);
                o0 = o0 && e0 <= b0; b0 = o0 ? e0 : b0;
                o1 = o1 && e1 <= b1; b1 = o1 ? e1 : b1;
            }
            tot0 += b0; tot1 += b1;
        }
        acc += tot0 + tot1;
        p[0][0] = add2(p[0][0], pk(1.0f, 1.0f));
    }
    out[blockIdx.x * 128 + threadIdx.x] = acc;
}
int main(){
    const int grid = 148 * 4, nt = 128, iters = 2000;
    float *tex, *pal, *out;
    cudaMalloc(&tex, grid * nt * 48 * 4); cudaMalloc(&pal, 24 * 4); cudaMalloc(&out, grid * nt * 4);
    float* h = (float*)malloc(grid * nt * 48 * 4); for (int i = 0; i < grid * nt * 48; i++) h[i] = (float)((i * 2654435761u) >> 17);
    cudaMemcpy(tex, h, grid * nt * 48 * 4, cudaMemcpyHostToDevice);
    float hp[24]; for (int i = 0; i < 24; i++) hp[i] = 1000.0f * i; cudaMemcpy(pal, hp, 96, cudaMemcpyHostToDevice);
    cudaEvent_t e0,e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int mode = 0; mode < 2; mode++) for (int rep = 0; rep < 3; rep++) {
        cudaEventRecord(e0);
        if (mode == 0) k_scalar<<<grid, nt>>>(tex, pal, out, iters); else k_packed<<<grid, nt>>>(tex, pal, out, iters, 0ull);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        const double pairs = (double)grid * nt * iters * 16 * 8 * (mode ? 2 : 1);
        printf("%s: %.3f ms, %.1f G pairs/s (%s)\n", mode ? "packed-2-signs" : "scalar", ms, pairs / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
