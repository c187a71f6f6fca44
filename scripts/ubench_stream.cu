// Micro-benchmark: which read pattern streams an int64 array fastest on B200?  Informs the layout of the memory-bound
// token kernels (rmpad count pass).  Build + run on the GPU box:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/ubench_stream scripts/ubench_stream.cu && /tmp/ubench_stream
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define CK(x)                                                                      \
    do {                                                                           \
        cudaError_t e = (x);                                                       \
        if (e != cudaSuccess) {                                                    \
            printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__);      \
            return 1;                                                              \
        }                                                                          \
    } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// K1: warp chunk of 512 elements, 16 x 8-byte loads per lane in flight, grid-stride over chunks
template <int BPS>
__global__ void __launch_bounds__(256, BPS) k_warp8(const long long* __restrict__ x, unsigned n, unsigned* out) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const unsigned nch = n / 512, tw = gridDim.x * 8;
    unsigned acc = 0;
    for (unsigned ch = blockIdx.x * 8 + w; ch < nch; ch += tw) {
        long long m[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) m[j] = x[ch * 512u + j * 32 + lane];
#pragma unroll
        for (int j = 0; j < 16; ++j) acc += __popc(__ballot_sync(0xffffffffu, m[j] != 0));
    }
    if (lane == 0 && acc == 0xdeadbeef) out[0] = acc;
}

// K2: warp chunk of 512 elements, 8 x 16-byte loads per lane
template <int BPS>
__global__ void __launch_bounds__(256, BPS) k_warp16(const long long* __restrict__ x, unsigned n, unsigned* out) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const unsigned nch = n / 512, tw = gridDim.x * 8;
    unsigned acc = 0;
    for (unsigned ch = blockIdx.x * 8 + w; ch < nch; ch += tw) {
        longlong2 m[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) m[j] = *reinterpret_cast<const longlong2*>(x + ch * 512u + j * 64 + lane * 2);
#pragma unroll
        for (int j = 0; j < 8; ++j)
            acc += __popc(__ballot_sync(0xffffffffu, m[j].x != 0)) + __popc(__ballot_sync(0xffffffffu, m[j].y != 0));
    }
    if (lane == 0 && acc == 0xdeadbeef) out[0] = acc;
}

// K3: flat grid-stride, VEC elements per thread per iteration (16-byte loads), high occupancy
template <int LOADS>
__global__ void __launch_bounds__(256) k_flat(const long long* __restrict__ x, unsigned n, unsigned* out) {
    unsigned acc = 0;
    const unsigned per_it = LOADS * 2;
    const unsigned stride = gridDim.x * 256u * per_it;
    for (unsigned i = (blockIdx.x * 256u + threadIdx.x) * 2; i < n; i += stride) {
        longlong2 m[LOADS];
#pragma unroll
        for (int j = 0; j < LOADS; ++j) {
            const unsigned k = i + j * gridDim.x * 512u;
            m[j] = (k < n) ? *reinterpret_cast<const longlong2*>(x + k) : make_longlong2(0, 0);
        }
#pragma unroll
        for (int j = 0; j < LOADS; ++j) acc += (m[j].x != 0) + (m[j].y != 0);
    }
    if (acc == 0xdeadbeef) out[0] = acc;
}


// K5: K1 + the side traffic of the rmpad count pass, switched on one by one (FLAGS bit 0: 64-byte bitmap store per
// chunk, bit 1: one RED per chunk on a per-tile counter, bit 2: __threadfence + ticket atomic per block at the end)
template <int FLAGS>
__global__ void __launch_bounds__(256, 4) k_side(const long long* __restrict__ x, unsigned n, unsigned* out,
                                                 unsigned* bits, int* tile_counts, unsigned* done) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const unsigned nch = n / 512, tw = gridDim.x * 8;
    unsigned acc = 0;
    for (unsigned ch = blockIdx.x * 8 + w; ch < nch; ch += tw) {
        long long m[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) m[j] = x[ch * 512u + j * 32 + lane];
        unsigned mine = 0;
        int cnt = 0;
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            const unsigned b = __ballot_sync(0xffffffffu, m[j] != 0);
            if (lane == j) mine = b;
            cnt += __popc(b);
        }
        acc += cnt;
        if ((FLAGS & 1) && lane < 16) bits[(size_t)ch * 16 + lane] = mine;
        if ((FLAGS & 2) && lane == 0) atomicAdd(&tile_counts[ch >> 3], cnt);
    }
    if (FLAGS & 4) {
        __threadfence();
        __syncthreads();
        if (threadIdx.x == 0) atomicAdd(done, 1u);
    }
    if (lane == 0 && acc == 0xdeadbeef) out[0] = acc;
}

// K4: warp chunk staged by ONE cp.async.bulk (4 KB) per chunk, STAGES in flight per warp
template <int STAGES>
__global__ void __launch_bounds__(256) k_tma(const long long* __restrict__ x, unsigned n, unsigned* out) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) uint64_t bars[8 * STAGES];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const unsigned nch = n / 512, tw = gridDim.x * 8, first = blockIdx.x * 8 + w;
    long long* st = reinterpret_cast<long long*>(smem) + (size_t)w * STAGES * 512;
    uint64_t* mb = bars + w * STAGES;
    auto issue = [&](unsigned ch, int s) {
        if (ch < nch) {
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&mb[s])), "r"(4096u)
                         : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                             smem_u32(st + s * 512)),
                         "l"(x + (size_t)ch * 512), "r"(4096u), "r"(smem_u32(&mb[s]))
                         : "memory");
        }
    };
    if (lane == 0) {
        for (int s = 0; s < STAGES; ++s)
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&mb[s])) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        for (int s = 0; s < STAGES; ++s) issue(first + s * tw, s);
    }
    __syncwarp();
    unsigned acc = 0, k = 0;
    for (unsigned ch = first; ch < nch; ch += tw, ++k) {
        const int s = k % STAGES;
        const unsigned parity = (k / STAGES) & 1;
        uint32_t ok = 0;
        while (!ok) {
            asm volatile(
                "{\n\t.reg .pred P;\n\tmbarrier.try_wait.parity.shared::cta.b64 P, [%1], %2;\n\tselp.u32 %0, 1, 0, P;\n\t}\n"
                : "=r"(ok)
                : "r"(smem_u32(&mb[s])), "r"(parity)
                : "memory");
        }
        const long long* buf = st + s * 512;
#pragma unroll
        for (int j = 0; j < 16; ++j) acc += __popc(__ballot_sync(0xffffffffu, buf[j * 32 + lane] != 0));
        __syncwarp();
        if (lane == 0) issue(ch + STAGES * tw, s);
    }
    if (lane == 0 && acc == 0xdeadbeef) out[0] = acc;
}

template <typename F>
static float time_it(F launch, int reps) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    for (int i = 0; i < 3; ++i) launch();
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    for (int i = 0; i < reps; ++i) launch();
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    return ms / reps;
}

int main() {
    const unsigned n = 1u << 25;
    long long* x;
    unsigned* out;
    CK(cudaMalloc(&x, (size_t)n * 8));
    CK(cudaMalloc(&out, 16));
    CK(cudaMemset(x, 1, (size_t)n * 8));
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    const double gb = (double)n * 8 / 1e9;
    auto report = [&](const char* name, float ms) { printf("%-44s %8.1f us %7.0f GB/s\n", name, ms * 1e3, gb / (ms * 1e-3)); };
    report("warp8  16x8B/lane, 4 blocks/SM", time_it([&] { k_warp8<4><<<sms * 4, 256>>>(x, n, out); }, 20));
    report("warp8  16x8B/lane, 2 blocks/SM", time_it([&] { k_warp8<2><<<sms * 2, 256>>>(x, n, out); }, 20));
    report("warp16 8x16B/lane, 4 blocks/SM", time_it([&] { k_warp16<4><<<sms * 4, 256>>>(x, n, out); }, 20));
    report("warp16 8x16B/lane, 8192 blocks", time_it([&] { k_warp16<4><<<8192, 256>>>(x, n, out); }, 20));
    report("flat   1x16B/thread, 8 blocks/SM", time_it([&] { k_flat<1><<<sms * 8, 256>>>(x, n, out); }, 20));
    report("flat   2x16B/thread, 8 blocks/SM", time_it([&] { k_flat<2><<<sms * 8, 256>>>(x, n, out); }, 20));
    report("flat   4x16B/thread, 8 blocks/SM", time_it([&] { k_flat<4><<<sms * 8, 256>>>(x, n, out); }, 20));
    report("flat   8x16B/thread, 4 blocks/SM", time_it([&] { k_flat<8><<<sms * 4, 256>>>(x, n, out); }, 20));
    CK(cudaFuncSetAttribute(k_tma<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 4 * 4096));
    report("tma    4 KB bulk x4 stages, 1 block/SM", time_it([&] { k_tma<4><<<sms, 256, 8 * 4 * 4096>>>(x, n, out); }, 20));
    CK(cudaFuncSetAttribute(k_tma<6>, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 6 * 4096));
    report("tma    4 KB bulk x6 stages, 1 block/SM", time_it([&] { k_tma<6><<<sms, 256, 8 * 6 * 4096>>>(x, n, out); }, 20));
    CK(cudaFuncSetAttribute(k_tma<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 2 * 4096));
    report("tma    4 KB bulk x2 stages, 3 blocks/SM", time_it([&] { k_tma<2><<<sms * 3, 256, 8 * 2 * 4096>>>(x, n, out); }, 20));
    unsigned* bits;
    int* tc;
    CK(cudaMalloc(&bits, (size_t)n / 8));
    CK(cudaMalloc(&tc, (size_t)(n / 4096) * 4 + 16));
    CK(cudaMemset(tc, 0, (size_t)(n / 4096) * 4 + 16));
    unsigned* done = reinterpret_cast<unsigned*>(tc + n / 4096);
    report("side: none", time_it([&] { k_side<0><<<sms * 4, 256>>>(x, n, out, bits, tc, done); }, 20));
    report("side: +bitmap store", time_it([&] { k_side<1><<<sms * 4, 256>>>(x, n, out, bits, tc, done); }, 20));
    report("side: +tile RED", time_it([&] { k_side<2><<<sms * 4, 256>>>(x, n, out, bits, tc, done); }, 20));
    report("side: +bitmap +RED", time_it([&] { k_side<3><<<sms * 4, 256>>>(x, n, out, bits, tc, done); }, 20));
    report("side: +bitmap +RED +fence", time_it([&] { k_side<7><<<sms * 4, 256>>>(x, n, out, bits, tc, done); }, 20));
    CK(cudaDeviceSynchronize());
    CK(cudaGetLastError());
    return 0;
}
