// tools/bench_radix.cu -- developer microbenchmark of the hand-written onesweep radix sort
// (pfbwt-f_b200/csrc/radix.cuh) against cub::DeviceRadixSort on the dictionary-sort shape:
// n (u64 key, u32 value) pairs, `bits` significant key bits starting at bit 10.
//   nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo --extended-lambda -I pfbwt-f_b200/csrc -I include -o bin/bench_radix tools/bench_radix.cu
#include "prim.cuh"
namespace pfb {
size_t &dev_bytes_live() { static size_t v = 0; return v; }
size_t &dev_bytes_peak() { static size_t v = 0; return v; }
SlabPool &slab_pool() { static SlabPool *p = new SlabPool(); return *p; }
}
using namespace pfb;
__host__ __device__ inline u64 mix(u64 x) { x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull; x ^= x >> 27; x *= 0x94D049BB133111EBull; x ^= x >> 31; return x; }
template <class C>
static void run_ours(const char *name, u64 n, int bits, int reps, DevBuf<u64> &ka, DevBuf<u64> &kb, DevBuf<u32> &va, DevBuf<u32> &vb,
                     Scratch &sc, cudaStream_t st) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    DevBuf<u32> bad(2);
    float best = 1e30f; u32 hb[2] = {0, 0};
    for (int r = 0; r < reps; ++r) {
        u64 *k = ka.get(); const u64 msk = (bits >= 54 ? ~0ull : ((1ull << bits) - 1)) << 10;
        launch_for(n, st, [=] __device__(u64 i) { k[i] = (mix(i * 0x9E3779B97F4A7C15ull + 7) & msk) | (i & 1023); });
        cudaEventRecord(e0, st);
        radix_sort_pairs<C>(ka, kb, va, vb, n, 10, 10 + bits, true, nullptr, sc.rw, st);
        cudaEventRecord(e1, st); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
        cudaMemsetAsync(bad.get(), 0, 8, st);
        const u64 *ks = ka.get(); const u32 *vs = va.get(); u32 *bp = bad.get();
        launch_for(n, st, [=] __device__(u64 i) {
            u64 a = ks[i] >> 10;
            u64 expect = (mix((u64)vs[i] * 0x9E3779B97F4A7C15ull + 7) & msk) | ((u64)vs[i] & 1023);
            if (expect != ks[i]) atomicAdd(bp + 1, 1u);
            if (i) { u64 b = ks[i - 1] >> 10; if (b > a || (b == a && vs[i - 1] >= vs[i])) atomicAdd(bp, 1u); }
        });
        u32 h2[2]; cudaMemcpyAsync(h2, bad.get(), 8, cudaMemcpyDeviceToHost, st); cudaStreamSynchronize(st);
        hb[0] += h2[0]; hb[1] += h2[1];
    }
    printf("%-22s n=%llu bits=%d: best %.3f ms  errors=%u/%u\n", name, (unsigned long long)n, bits, best, hb[0], hb[1]);
    fflush(stdout);
}
#define RUN(NT, IPT, MINB, MATCH) run_ours<RxCfg<NT, IPT, MINB, MATCH>>("ours " #NT "x" #IPT " b" #MINB " m" #MATCH, n, bits, reps, ka, kb, va, vb, sc, st)

int main(int argc, char **argv) {
    u64 n = argc > 1 ? strtoull(argv[1], 0, 10) : 1100000000ull;
    int bits = argc > 2 ? atoi(argv[2]) : 40;
    int reps = argc > 3 ? atoi(argv[3]) : 3;
    cudaStream_t st; cudaStreamCreate(&st);
    alloc_stream() = st;
    Scratch sc;
    DevBuf<u64> ka(n), kb(n); DevBuf<u32> va(n), vb(n);
    DevBuf<u32> bad(2);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    if (argc > 4) {
        RUN(256, 16, 3, 0); RUN(256, 20, 2, 0); RUN(256, 20, 3, 0); RUN(256, 24, 2, 0); RUN(256, 28, 2, 0); RUN(256, 32, 1, 0);
        RUN(384, 16, 1, 0); RUN(384, 20, 1, 0); RUN(512, 16, 1, 0); RUN(256, 12, 4, 0);
    }
    for (int impl = 0; impl < 2; ++impl) {
        for (int r = 0; r < reps; ++r) {
            u64 *k = ka.get(); const u64 msk = (bits >= 54 ? ~0ull : ((1ull << bits) - 1)) << 10;
            launch_for(n, st, [=] __device__(u64 i) { k[i] = (mix(i * 0x9E3779B97F4A7C15ull + 7) & msk) | (i & 1023); });
            cudaEventRecord(e0, st);
            if (impl == 0) radix_sort_pairs(ka, kb, va, vb, n, 10, 10 + bits, true, nullptr, sc.rw, st);
            else {
                u32 *v = va.get(); launch_for(n, st, [=] __device__(u64 i) { v[i] = (u32)i; });
                cub::DoubleBuffer<u64> dk(ka.get(), kb.get()); cub::DoubleBuffer<u32> dv(va.get(), vb.get());
                size_t bytes = 0; cub::DeviceRadixSort::SortPairs(nullptr, bytes, dk, dv, (int64_t)n, 10, 10 + bits, st);
                void *tmp = sc.ensure(bytes); cub::DeviceRadixSort::SortPairs(tmp, bytes, dk, dv, (int64_t)n, 10, 10 + bits, st);
                if (dk.Current() != ka.get()) ka.swap(kb);
                if (dv.Current() != va.get()) va.swap(vb);
            }
            cudaEventRecord(e1, st); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            // check: sorted on the bit range, stable (values ascending inside equal keys), values consistent with keys
            cudaMemsetAsync(bad.get(), 0, 8, st);
            const u64 *ks = ka.get(); const u32 *vs = va.get(); u32 *bp = bad.get();
            launch_for(n, st, [=] __device__(u64 i) {
                u64 a = ks[i] >> 10;
                u64 expect = (mix((u64)vs[i] * 0x9E3779B97F4A7C15ull + 7) & msk) | ((u64)vs[i] & 1023);
                if (expect != ks[i]) atomicAdd(bp + 1, 1u);
                if (i) { u64 b = ks[i - 1] >> 10; if (b > a || (b == a && vs[i - 1] >= vs[i])) atomicAdd(bp, 1u); }
            });
            u32 hb[2]; cudaMemcpyAsync(hb, bad.get(), 8, cudaMemcpyDeviceToHost, st); cudaStreamSynchronize(st);
            printf("%s n=%llu bits=%d: %.3f ms  (%.1f GB/s of 24 B/pair/pass)  order_errors=%u payload_errors=%u\n", impl ? "cub " : "ours",
                   (unsigned long long)n, bits, ms, (double)n * 24.0 * ((bits + 7) / 8) / ms / 1e6, hb[0], hb[1]);
        }
    }
    return 0;
}
