// Variant 1 of the copy+checksum kernel: TMA bulk staging.
//
// Same contract as copy_csum_kernel (dissem_kernels.cu) — same descriptor table, same work units, same
// checksum, same release of the destination's received counter — but the bytes travel
//   global --cp.async.bulk--> shared (mbarrier ring) --cp.async.bulk--> global (local HBM or a peer GPU)
// and only the checksum touches registers (LDS.128 from the staged tile).  Warp roles per CTA:
//   warp 0, lane 0   producer: draws tickets, cuts a unit's 16-byte-aligned body into stages, issues the loads
//   warp 1, lane 0   storer:   issues the bulk stores, retires stages, adds the unit's (A, B) and received bytes
//   warps 2..5       consumers: checksum of every staged tile (and the multimem.st of a multicast unit)
// Unaligned head/tail bytes of a fragment (at most 15 + 15 per unit) are moved by the storer with plain ld/st.
#include "dissem_kernels.cuh"

namespace dissem {

namespace {

// Ring shape is a template parameter: kStageBytes per bulk copy, kStages slots, kStoreLag bulk-store groups allowed in
// flight before a stage is retired (needs kStoreLag + 2 <= kStages).
// kCW consumer warps per CTA is a template parameter too: 4 are enough to checksum tiles at HBM speed; multicast units,
// whose stores are issued by the consumers (multimem.st from registers), get wider CTAs (see launch_copy_csum_tma).
constexpr int kMaxConsumerWarps = 14;

enum : int { kMetaLast = 1, kMetaEnd = 2 };

struct StageMeta {
    const uint8_t *src;
    uint8_t *dst;
    unsigned long long *csum;
    unsigned long long *received;
    long long pos;        // layer offset of this stage's first byte (16-byte aligned)
    long long b0, b1;     // the unit's byte range in layer space
    int n;                // bytes staged (multiple of 16, may be 0)
    int flags;            // kMetaLast: last stage of its unit; kMetaEnd: no more work
    int mc;               // bit 0: multicast unit; bit 1: whole-chunk unit whose sums are stored (kFlagStoreSums)
    int pad;
};

template <int kStageBytes, int kStages>
struct __align__(128) SmemLayout {
    unsigned char data[kStages][kStageBytes];
    StageMeta meta[kStages];
    unsigned long long usum[kStages][2];
    unsigned long long red[2][2][kMaxConsumerWarps];
    unsigned long long full[kStages];
    unsigned long long empty[kStages];
    unsigned long long cons_done[kStages];
};

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(unsigned long long *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(unsigned long long *bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WAIT_DONE;\n"
        "bra WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void tma_load(void *smem_dst, const void *gmem_src, unsigned bytes, unsigned long long *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void tma_store(void *gmem_dst, const void *smem_src, unsigned bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst), "r"(smem_u32(smem_src)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void tma_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void tma_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void tma_wait_all() {
    asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void st_multicast16(void *p, const uint4 &v) {
    asm volatile("multimem.st.weak.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(__uint_as_float(v.x)),
                 "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)), "f"(__uint_as_float(v.w))
                 : "memory");
}
__device__ __forceinline__ void st_multicast64(unsigned long long *p, unsigned long long v) {
    asm volatile("multimem.st.relaxed.sys.global.b64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void red_multicast64(unsigned long long *p, unsigned long long v) {
    asm volatile("multimem.red.relaxed.sys.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long warp_sum64(unsigned long long v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

template <int kStageBytes, int kStages, int kStoreLag, int kConsumerWarps = 4>
__global__ void __launch_bounds__(64 + 32 * kConsumerWarps, 2)
copy_csum_tma_kernel(const XferDesc *__restrict__ descs, int n_descs, long long total_units, int unit_shift,
                     SchedCounters *counters) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    static_assert(kStoreLag + 2 <= kStages, "ring too shallow for the store lag");
    static_assert(kConsumerWarps <= kMaxConsumerWarps, "red[] too small");
    using Smem = SmemLayout<kStageBytes, kStages>;
    Smem &sm = *reinterpret_cast<Smem *>(smem_raw);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    if (tid == 0) {
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&sm.full[s], 1);
            mbar_init(&sm.empty[s], 1);
            mbar_init(&sm.cons_done[s], kConsumerWarps);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    if (warp == 0) {
        // ------------------------------------------------------------------ producer
        if (lane == 0) {
            long long i = 0;                                     // running stage count
            for (;;) {
                const long long u = (long long)atomicAdd(&counters->ticket, 1ull);
                const bool end = u >= total_units;
                const XferDesc *D = nullptr;
                long long b0 = 0, b1 = 0, a0 = 0, a1 = 0;
                if (!end) {
                    long long slice_idx;
                    D = descs + locate_unit(descs, n_descs, u, &slice_idx);
                    const long long off = D->off, size = D->size;
                    const long long slice = (off >> unit_shift) + slice_idx;
                    b0 = slice << unit_shift;
                    b1 = b0 + (1ll << unit_shift);
                    if (b0 < off) b0 = off;
                    if (b1 > off + size) b1 = off + size;
                    a0 = (b0 + 15) & ~15ll;
                    a1 = b1 & ~15ll;
                    if (a0 > b1) a0 = b1;
                    if (a1 < a0) a1 = a0;
                    // relay hop: the source copy is still arriving; every read of this unit (the bulk loads below, the
                    // storer's head/tail bytes) happens after this point
                    if (D->wait_ptr) wait_for_source(D->wait_ptr, D->wait_value, counters);
                }
                long long pos = a0;
                do {
                    const int s = (int)(i % kStages);
                    const unsigned par = (unsigned)((i / kStages) & 1);
                    mbar_wait(&sm.empty[s], par ^ 1u);           // slot retired by the storer (first lap passes)
                    StageMeta &m = sm.meta[s];
                    if (end) {
                        m.flags = kMetaEnd;
                        m.n = 0;
                        mbar_arrive(&sm.full[s]);
                    } else {
                        long long n = a1 - pos;
                        if (n > kStageBytes) n = kStageBytes;
                        m.src = D->src;
                        m.dst = D->dst;
                        m.csum = D->csum;
                        m.received = D->received;
                        m.pos = pos;
                        m.b0 = b0;
                        m.b1 = b1;
                        m.n = (int)n;
                        {
                            // bit 0: multicast unit; bit 1: the unit covers its whole chunk and the sums may be stored
                            long long ce = ((b0 >> kChunkShift) + 1) << kChunkShift;
                            if (ce > D->total) ce = D->total;
                            const bool whole = (D->flags & kFlagStoreSums) && (b0 & (kChunkBytes - 1)) == 0 && b1 == ce;
                            m.mc = ((D->flags & kFlagMulticast) ? 1 : 0) | (whole ? 2 : 0);
                        }
                        m.flags = (pos + n >= a1) ? kMetaLast : 0;
                        if (n > 0) {
                            mbar_arrive_expect_tx(&sm.full[s], (unsigned)n);
                            tma_load(sm.data[s], D->src + pos, (unsigned)n, &sm.full[s]);
                        } else {
                            mbar_arrive(&sm.full[s]);
                        }
                        pos += n;
                    }
                    ++i;
                } while (!end && pos < a1);
                if (end) break;
            }
            // re-arm the scheduling counters for the next launch (last CTA out)
            __threadfence();
            const unsigned int prev = atomicAdd(&counters->done_ctas, 1u);
            if (prev == gridDim.x - 1) {
                counters->ticket = 0ull;
                counters->done_ctas = 0u;
                __threadfence();
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ storer
        if (lane == 0) {
            long long i = 0;          // stage being issued
            long long r = 0;          // stage being retired (r <= i)
            bool done = false;
            auto retire = [&](long long j, bool drain) {
                const int s = (int)(j % kStages);
                const unsigned par = (unsigned)((j / kStages) & 1);
                const StageMeta m = sm.meta[s];
                mbar_wait(&sm.cons_done[s], par);                 // consumers are done with the tile (and usum is written)
                if (m.flags & kMetaLast) {
                    // every bulk store of this unit has been issued (stages retire in order); wait for their completion
                    if (drain) tma_wait_all<0>(); else tma_wait_all<kStoreLag>();
                    unsigned long long A = sm.usum[s][0], S2 = sm.usum[s][1];
                    const long long base_word = m.b0 >> 2;
                    long long a0 = (m.b0 + 15) & ~15ll, a1 = m.b1 & ~15ll;
                    if (a0 > m.b1) a0 = m.b1;
                    if (a1 < a0) a1 = a0;
                    for (long long p = m.b0; p < m.b1; ++p) {     // head [b0,a0) and tail [a1,b1)
                        if (p == a0) { p = a1; if (p >= m.b1) break; }
                        const unsigned int b = m.src[p];
                        if (m.dst) m.dst[p] = (uint8_t)b;
                        const unsigned long long val = (unsigned long long)(b << (8 * (unsigned)(p & 3)));
                        A += val;
                        S2 += (unsigned long long)((p >> 2) - base_word) * val;
                    }
                    const unsigned long long B = (unsigned long long)(base_word + 1) * A + S2;
                    if (m.csum) {
                        const long long c = m.b0 >> kChunkShift;
                        if (m.mc & 2) {              // the unit is its chunk (kFlagStoreSums): stored, no prior zeroing needed
                            if (m.mc & 1) { st_multicast64(m.csum + 2 * c, A); st_multicast64(m.csum + 2 * c + 1, B); }
                            else { m.csum[2 * c] = A; m.csum[2 * c + 1] = B; }
                        } else if (m.mc & 1) { red_multicast64(m.csum + 2 * c, A); red_multicast64(m.csum + 2 * c + 1, B); }
                        else { atomicAdd(m.csum + 2 * c, A); atomicAdd(m.csum + 2 * c + 1, B); }
                    }
                    if (m.received) {
                        __threadfence_system();
                        if (m.mc & 1) red_multicast64(m.received, (unsigned long long)(m.b1 - m.b0));
                        else atomicAdd_system(m.received, (unsigned long long)(m.b1 - m.b0));
                    }
                } else {
                    if (drain) tma_wait_read<0>(); else tma_wait_read<kStoreLag>();   // the tile has been read out
                }
                mbar_arrive(&sm.empty[s]);
            };
            while (!done) {
                const int s = (int)(i % kStages);
                const unsigned par = (unsigned)((i / kStages) & 1);
                mbar_wait(&sm.full[s], par);
                const int flags = sm.meta[s].flags;
                if (flags & kMetaEnd) {
                    done = true;
                } else {
                    const int n = sm.meta[s].n;
                    uint8_t *dst = sm.meta[s].dst;
                    if (dst && !(sm.meta[s].mc & 1) && n > 0) {
                        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                        tma_store(dst + sm.meta[s].pos, sm.data[s], (unsigned)n);
                    }
                    tma_commit();                                  // one group per stage, empty or not
                    ++i;
                    if (i - r > kStoreLag) { retire(r, false); ++r; }
                }
            }
            while (r < i) { retire(r, true); ++r; }
        }
    } else {
        // ------------------------------------------------------------------ consumers
        const int cw = warp - 2;
        const int ctid = tid - 64;
        constexpr int kCThreads = 32 * kConsumerWarps;
        unsigned long long a = 0ull, s2 = 0ull;
        int redp = 0;
        for (long long i = 0;; ++i) {
            const int s = (int)(i % kStages);
            const unsigned par = (unsigned)((i / kStages) & 1);
            mbar_wait(&sm.full[s], par);
            const int flags = sm.meta[s].flags;
            if (flags & kMetaEnd) break;
            const int n = sm.meta[s].n;
            const long long pos = sm.meta[s].pos;
            const long long base_word = sm.meta[s].b0 >> 2;
            const unsigned int k_stage = (unsigned int)((pos >> 2) - base_word);
            const uint4 *tile = reinterpret_cast<const uint4 *>(sm.data[s]);
            const int nvec = n >> 4;
            const bool mc = (sm.meta[s].mc & 1) != 0 && sm.meta[s].dst != nullptr;
            uint4 *mcdst = mc ? reinterpret_cast<uint4 *>(sm.meta[s].dst + pos) : nullptr;
            int v = ctid;
            if (mc) {
                // multicast unit: the stores leave from registers.  Four tiles are read, their four multimem.st issued back
                // to back, and only then folded into the checksum, so that a thread keeps several stores in flight.
                for (; v + 3 * kCThreads < nvec; v += 4 * kCThreads) {
                    uint4 w[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) w[j] = tile[v + j * kCThreads];
#pragma unroll
                    for (int j = 0; j < 4; ++j) st_multicast16(mcdst + v + j * kCThreads, w[j]);
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const unsigned int k0 = k_stage + 4u * (unsigned int)(v + j * kCThreads);
                        a += (unsigned long long)w[j].x + w[j].y + w[j].z + w[j].w;
                        s2 += (unsigned long long)k0 * w[j].x + (unsigned long long)(k0 + 1) * w[j].y +
                              (unsigned long long)(k0 + 2) * w[j].z + (unsigned long long)(k0 + 3) * w[j].w;
                    }
                }
            }
            for (; v < nvec; v += kCThreads) {
                const uint4 w = tile[v];
                if (mc) st_multicast16(mcdst + v, w);
                const unsigned int k0 = k_stage + 4u * (unsigned int)v;
                a += (unsigned long long)w.x + w.y + w.z + w.w;
                s2 += (unsigned long long)k0 * w.x + (unsigned long long)(k0 + 1) * w.y +
                      (unsigned long long)(k0 + 2) * w.z + (unsigned long long)(k0 + 3) * w.w;
            }
            if (flags & kMetaLast) {
                // unit complete: reduce the consumers' partial sums and hand (A, S2) to the storer through usum[s]
                const unsigned long long wa = warp_sum64(a), ws = warp_sum64(s2);
                if (lane == 0) { sm.red[redp][0][cw] = wa; sm.red[redp][1][cw] = ws; }
                asm volatile("bar.sync 1, %0;" ::"n"(kCThreads) : "memory");
                if (ctid == 0) {
                    unsigned long long A = 0ull, S2 = 0ull;
#pragma unroll
                    for (int w2 = 0; w2 < kConsumerWarps; ++w2) { A += sm.red[redp][0][w2]; S2 += sm.red[redp][1][w2]; }
                    sm.usum[s][0] = A;
                    sm.usum[s][1] = S2;
                }
                redp ^= 1;          // red[] is double-buffered: its next use is two units away, behind another bar.sync
                a = 0ull;
                s2 = 0ull;
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm.cons_done[s]);        // after usum is written by this warp's lane 0 (warp 2)
        }
    }
}

template <int kStageBytes, int kStages, int kStoreLag, int kConsumerWarps = 4>
static cudaError_t launch_tma(const XferDesc *descs, int n_descs, long long total_units, int unit_shift,
                              SchedCounters *counters, const LaunchCfg &cfg, cudaStream_t stream, int *grid_out) {
    using Smem = SmemLayout<kStageBytes, kStages>;
    static_assert(sizeof(Smem) <= 112 * 1024, "two CTAs per SM");
    auto kern = copy_csum_tma_kernel<kStageBytes, kStages, kStoreLag, kConsumerWarps>;
    constexpr int kThreadsTma = 64 + 32 * kConsumerWarps;
    const int smem = (int)sizeof(Smem) + 128;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    int per_sm = cfg.ctas_per_sm;
    if (per_sm <= 0) {
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kThreadsTma, smem);
        if (e != cudaSuccess) return e;
        if (per_sm < 1) per_sm = 1;
    }
    long long grid = (long long)cfg.n_sm * per_sm;
    if (grid > total_units) grid = total_units;
    if (grid < 1) grid = 1;
    if (grid_out) *grid_out = (int)grid;
    kern<<<(int)grid, kThreadsTma, smem, stream>>>(descs, n_descs, total_units, unit_shift, counters);
    return cudaGetLastError();
}

}  // namespace

cudaError_t warm_kernels_tma() {
    cudaFuncAttributes a;
    return cudaFuncGetAttributes(&a, copy_csum_tma_kernel<16384, 6, 3>);
}

cudaError_t launch_copy_csum_tma(const XferDesc *descs, int n_descs, long long total_units, int unit_shift,
                                 SchedCounters *counters, const LaunchCfg &cfg, cudaStream_t stream, int *grid_out) {
    switch (cfg.tma_cfg) {
        case 1: return launch_tma<32768, 3, 1>(descs, n_descs, total_units, unit_shift, counters, cfg, stream, grid_out);
        case 2: return launch_tma<8192, 12, 6>(descs, n_descs, total_units, unit_shift, counters, cfg, stream, grid_out);
        case 3: return launch_tma<16384, 4, 2>(descs, n_descs, total_units, unit_shift, counters, cfg, stream, grid_out);
        case 4: return launch_tma<8192, 6, 3>(descs, n_descs, total_units, unit_shift, counters, cfg, stream, grid_out);
        case 5: return launch_tma<4096, 12, 6>(descs, n_descs, total_units, unit_shift, counters, cfg, stream, grid_out);
        // wider CTAs (8 / 14 consumer warps): the shapes for multicast units, whose stores the consumers issue
        case 6: return launch_tma<16384, 6, 3, 8>(descs, n_descs, total_units, unit_shift, counters, cfg, stream, grid_out);
        case 7: return launch_tma<16384, 6, 3, 14>(descs, n_descs, total_units, unit_shift, counters, cfg, stream, grid_out);
        case 8: return launch_tma<8192, 6, 3, 14>(descs, n_descs, total_units, unit_shift, counters, cfg, stream, grid_out);
        default: return launch_tma<16384, 6, 3>(descs, n_descs, total_units, unit_shift, counters, cfg, stream, grid_out);
    }
}

}  // namespace dissem
