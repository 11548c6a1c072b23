// microbench_sp.cu — limits of the Schur-product access pattern on B200 (not part of the product):
//   1. streaming read of the E blocks (288 B per tag observation) with 16 B loads, many CTAs
//   2. the same stream through cp.async.bulk into a shared-memory ring (persistent CTAs)
//   3. RED.f64 scatter: 6 doubles per tag observation to its camera's slot (marker order)
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o microbench_sp microbench_sp.cu
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__global__ void __launch_bounds__(256) k_stream(const double2* __restrict__ E, size_t n16, double* out) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    double s = 0.0;
    for (; i + 3 * stride < n16; i += 4 * stride) {
        double2 a = __ldcs(E + i), b = __ldcs(E + i + stride), c = __ldcs(E + i + 2 * stride), d = __ldcs(E + i + 3 * stride);
        s += a.x + a.y + b.x + b.y + c.x + c.y + d.x + d.y;
    }
    for (; i < n16; i += stride) { double2 a = __ldcs(E + i); s += a.x + a.y; }
    if (s == 1.2345e-300) out[0] = s;
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* b) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
    asm volatile(
        "{\n.reg .pred p;\nWAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}" ::"r"(smem_u32(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// persistent: CTA b streams its contiguous byte range through a ring of NST stages of CH bytes
template <int NST, int CH>
__global__ void __launch_bounds__(256) k_bulk(const char* __restrict__ E, size_t bytes, double* out) {
    extern __shared__ __align__(128) char smem[];
    __shared__ uint64_t full[NST], empty[NST];
    const size_t nch = bytes / CH;
    const size_t c0 = nch * blockIdx.x / gridDim.x, c1 = nch * (blockIdx.x + 1) / gridDim.x;
    if (threadIdx.x == 0) {
        for (int s = 0; s < NST; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 8); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const int n = (int)(c1 - c0);
    if (threadIdx.x == 0) {
        for (int s = 0; s < NST && s < n; ++s) {
            mbar_expect_tx(&full[s], CH);
            bulk_g2s(smem + (size_t)s * CH, E + (c0 + s) * CH, CH, &full[s]);
        }
    }
    double acc = 0.0;
    for (int i = 0; i < n; ++i) {
        const int s = i % NST;
        const uint32_t ph = (i / NST) & 1;
        mbar_wait(&full[s], ph);
        const double2* p = reinterpret_cast<const double2*>(smem + (size_t)s * CH);
        for (int t = threadIdx.x; t < CH / 16; t += 256) { double2 v = p[t]; acc += v.x + v.y; }
        __syncwarp();
        if ((threadIdx.x & 31) == 0) mbar_arrive(&empty[s]);
        if (threadIdx.x == 0 && i + NST < n) {
            mbar_wait(&empty[s], ph);
            mbar_expect_tx(&full[s], CH);
            bulk_g2s(smem + (size_t)s * CH, E + (c0 + i + NST) * CH, CH, &full[s]);
        }
    }
    if (acc == 1.2345e-300) out[0] = acc;
}

// RED scatter in marker order: thread per tag observation, 6 REDs to out[cam*6 + a]
__global__ void __launch_bounds__(256) k_red(const int* __restrict__ cam, int n, double* __restrict__ out) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const int c = cam[k];
    double* o = out + (size_t)c * 6;
#pragma unroll
    for (int a = 0; a < 6; ++a) atomicAdd(o + a, 1.0 + a);
}
// same with the camera vector padded to 8 doubles (64 B aligned slots: every block in 2 whole sectors)
__global__ void __launch_bounds__(256) k_red8(const int* __restrict__ cam, int n, double* __restrict__ out) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const int c = cam[k];
    double* o = out + (size_t)c * 8;
#pragma unroll
    for (int a = 0; a < 6; ++a) atomicAdd(o + a, 1.0 + a);
}
// gather only: 6 doubles per observation from x[cam]
__global__ void __launch_bounds__(256) k_gather(const int* __restrict__ cam, int n, const double* __restrict__ x, double* out) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const int c = cam[k];
    const double2* p = reinterpret_cast<const double2*>(x + (size_t)c * 6);
    double2 a = __ldg(p), b = __ldg(p + 1), d = __ldg(p + 2);
    const double s = a.x + a.y + b.x + b.y + d.x + d.y;
    if (s == 1.2345e-300) out[0] = s;
}

template <typename F>
float time_ms(F f, int reps = 20) {
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    for (int i = 0; i < 3; ++i) f();
    cudaEventRecord(a);
    for (int i = 0; i < reps; ++i) f();
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    return ms / reps;
}

int main() {
    const int Nm = 20000, per = 200, Nc = 100000;
    const int n = Nm * per;
    const size_t bytes = (size_t)n * 288;
    char* E; CK(cudaMalloc(&E, bytes)); CK(cudaMemset(E, 0, bytes));
    double* out; CK(cudaMalloc(&out, sizeof(double) * Nc * 8)); CK(cudaMemset(out, 0, sizeof(double) * Nc * 8));
    // cameras of marker j: 200 distinct cameras from a window of 1000 around 5*j (spatial locality)
    std::vector<int> cam(n);
    uint64_t st = 12345;
    for (int j = 0; j < Nm; ++j) {
        const int base = (int)((int64_t)j * (Nc - 1000) / Nm);
        for (int i = 0; i < per; ++i) {
            st = st * 6364136223846793005ull + 1442695040888963407ull;
            cam[(size_t)j * per + i] = base + (i * 5 + (int)((st >> 33) % 5));
        }
    }
    int* dcam; CK(cudaMalloc(&dcam, sizeof(int) * n)); CK(cudaMemcpy(dcam, cam.data(), sizeof(int) * n, cudaMemcpyHostToDevice));

    for (int g : { 148 * 2, 148 * 4, 148 * 8, 148 * 16, 148 * 32 }) {
        float ms = time_ms([&] { k_stream<<<g, 256>>>((const double2*)E, bytes / 16, out); });
        printf("stream LDG.128 grid %5d: %.4f ms  %.0f GB/s\n", g, ms, bytes / ms * 1e-6);
    }
    {
        constexpr int CH = 9216;
        auto run = [&](auto kern, int nst, int grid) {
            CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, nst * CH));
            float ms = time_ms([&] { kern<<<grid, 256, nst * CH>>>(E, bytes, out); });
            CK(cudaGetLastError());
            printf("bulk ring %2d x %d B, grid %4d: %.4f ms  %.0f GB/s\n", nst, CH, grid, ms, bytes / ms * 1e-6);
        };
        run(k_bulk<4, CH>, 4, 148); run(k_bulk<8, CH>, 8, 148); run(k_bulk<16, CH>, 16, 148); run(k_bulk<24, CH>, 24, 148);
        run(k_bulk<4, CH>, 4, 296); run(k_bulk<8, CH>, 8, 296); run(k_bulk<12, CH>, 12, 296);
        run(k_bulk<4, CH>, 4, 592); run(k_bulk<6, CH>, 6, 592);
    }
    {
        float ms = time_ms([&] { k_red<<<(n + 255) / 256, 256>>>(dcam, n, out); });
        printf("RED.f64 6 per obs (48 B slots): %.4f ms  %.1f G atomics/s\n", ms, 6.0 * n / ms * 1e-6);
        ms = time_ms([&] { k_red8<<<(n + 255) / 256, 256>>>(dcam, n, out); });
        printf("RED.f64 6 per obs (64 B slots): %.4f ms  %.1f G atomics/s\n", ms, 6.0 * n / ms * 1e-6);
        ms = time_ms([&] { k_gather<<<(n + 255) / 256, 256>>>(dcam, n, out + 0, out); });
        printf("gather 48 B per obs: %.4f ms\n", ms);
    }
    CK(cudaDeviceSynchronize());
    return 0;
}
