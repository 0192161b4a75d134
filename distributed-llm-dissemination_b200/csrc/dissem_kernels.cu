// sm_100a kernels of the layer-dissemination hot path.
//
// What they replace in the reference (paths under /root/reference/distributor):
//   * sender:   writeWithLimit's 256 KiB conn.Write loop over data[Offset:Offset+DataSize]
//               (transport.go:342-345, :407-424)
//   * receiver: make(LayerData, LayerSize) + io.ReadFull (transport.go:201-207) and the fragment
//               accounting of FlowRetransmitReceiverNode.handleLayerMsg (node.go:1542-1565)
// as ONE push kernel per source GPU: coalesced 16-byte loads from the source layer, 16-byte stores into
// the destination layer (local HBM, a peer GPU over NVLink, or an NVLS multicast address), with the
// per-chunk checksum (A = sum w, B = sum (i+1) w over little-endian u32 words, mod 2^64) accumulated
// from the registers that carry the data, and a release of the destination's received-bytes counter.
//
// The work is HBM/NVLink-bound byte moving: no tensor cores.  Persistent CTAs pull 256 KiB-or-smaller
// work units from a ticket counter (dynamic scheduling: local and remote units cost differently).
#include "dissem_kernels.cuh"

namespace dissem {

// ------------------------------------------------------------------------------------------------
// small PTX helpers
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 ld_stream(const uint4 *p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}
// A multicast unit rewrites the source's own copy through the multicast object while it is being read (the source is a
// member): not read-only for the kernel's lifetime, so no .nc path for it.
__device__ __forceinline__ uint4 ld_plain(const uint4 *p) {
    uint4 r;
    asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}
__device__ __forceinline__ void st_stream(uint4 *p, const uint4 &v) {
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y),
                 "r"(v.z), "r"(v.w)
                 : "memory");
}
// 256-bit forms (sm_100+): a warp then moves 1 KiB of contiguous bytes per instruction.
struct U8 { uint4 lo, hi; };
__device__ __forceinline__ U8 ld_stream32(const void *p) {
    U8 r;
    asm volatile("ld.global.nc.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r.lo.x), "=r"(r.lo.y), "=r"(r.lo.z), "=r"(r.lo.w), "=r"(r.hi.x), "=r"(r.hi.y), "=r"(r.hi.z), "=r"(r.hi.w)
                 : "l"(p));
    return r;
}
__device__ __forceinline__ void st_stream32(void *p, const U8 &v) {
    asm volatile("st.global.L1::no_allocate.v8.u32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(v.lo.x), "r"(v.lo.y),
                 "r"(v.lo.z), "r"(v.lo.w), "r"(v.hi.x), "r"(v.hi.y), "r"(v.hi.z), "r"(v.hi.w)
                 : "memory");
}
// NVLS: one store lands in every GPU bound to the multicast object.
__device__ __forceinline__ void st_multicast(uint4 *p, const uint4 &v) {
    asm volatile("multimem.st.weak.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p),
                 "f"(__uint_as_float(v.x)), "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)),
                 "f"(__uint_as_float(v.w))
                 : "memory");
}
__device__ __forceinline__ void red_multicast_u64(unsigned long long *p, unsigned long long v) {
    asm volatile("multimem.red.relaxed.sys.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void st_multicast_u64(unsigned long long *p, unsigned long long v) {
    asm volatile("multimem.st.relaxed.sys.global.b64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

__device__ __forceinline__ unsigned long long warp_sum(unsigned long long v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Per-thread checksum accumulators.  S2 uses the word offset k from the unit's first word (k < 2^17, so
// k*w is one 32x32->64 multiply); the unit's B is (first_word_index + 1) * A + S2.
struct Acc {
    unsigned long long a, s2;
    __device__ __forceinline__ void vec(const uint4 &v, unsigned int k0) {
        a += (unsigned long long)v.x + v.y + v.z + v.w;
        s2 += (unsigned long long)k0 * v.x + (unsigned long long)(k0 + 1) * v.y +
              (unsigned long long)(k0 + 2) * v.z + (unsigned long long)(k0 + 3) * v.w;
    }
    __device__ __forceinline__ void byte(unsigned int b, unsigned int k, unsigned int lane) {
        unsigned long long val = (unsigned long long)(b << (8 * lane));
        a += val;
        s2 += (unsigned long long)k * val;
    }
};

// ------------------------------------------------------------------------------------------------
// K1/K4/K5 (register path): copy + checksum, checksum only when dst == NULL.
// ------------------------------------------------------------------------------------------------
template <int kThreads, int kUnroll, bool kWide = false>
__global__ void __launch_bounds__(kThreads)
copy_csum_kernel(const XferDesc *__restrict__ descs, int n_descs, long long total_units, int unit_shift,
                 SchedCounters *counters) {
    constexpr int kWarps = kThreads / 32;
    __shared__ long long s_unit[2];
    __shared__ unsigned long long s_red[2][2][kWarps];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    if (tid == 0) s_unit[0] = (long long)atomicAdd(&counters->ticket, 1ull);
    __syncthreads();

    for (int p = 0;; p ^= 1) {
        const long long u = s_unit[p];
        if (u >= total_units) break;
        if (tid == 0) s_unit[p ^ 1] = (long long)atomicAdd(&counters->ticket, 1ull);   // prefetch the next ticket

        long long slice_idx;
        const int d = locate_unit(descs, n_descs, u, &slice_idx);
        const XferDesc *__restrict__ D = descs + d;
        const long long off = __ldg(&D->off), size = __ldg(&D->size);
        const long long slice = (off >> unit_shift) + slice_idx;
        long long b0 = slice << unit_shift, b1 = b0 + (1ll << unit_shift);      // byte range in LAYER space
        if (b0 < off) b0 = off;
        if (b1 > off + size) b1 = off + size;
        const uint8_t *__restrict__ src = D->src;
        uint8_t *__restrict__ dst = D->dst;
        const bool mc = (__ldg(&D->flags) & kFlagMulticast) != 0;
        const unsigned long long *wait_ptr = D->wait_ptr;            // relay hop: the source is still arriving
        if (wait_ptr) {                                               // uniform per CTA
            if (tid == 0) wait_for_source(wait_ptr, D->wait_value, counters);
            __syncthreads();
        }

        const long long base_word = b0 >> 2;                    // word holding the unit's first byte
        long long a0 = (b0 + 15) & ~15ll, a1 = b1 & ~15ll;      // 16-byte aligned body [a0, a1)
        if (a0 > b1) a0 = b1;
        if (a1 < a0) a1 = a0;

        Acc acc{0ull, 0ull};

        // head / tail bytes (fragments of a flow plan start and end at arbitrary byte offsets)
        {
            const int nh = (int)(a0 - b0), nt = (int)(b1 - a1);
            if (tid < nh + nt) {
                const long long pos = tid < nh ? b0 + tid : a1 + (tid - nh);
                const unsigned int b = src[pos];
                if (dst) dst[pos] = (uint8_t)b;
                acc.byte(b, (unsigned int)((pos >> 2) - base_word), (unsigned int)(pos & 3));
            }
        }

        const long long nvec = (a1 - a0) >> 4;
        const uint4 *__restrict__ s4 = reinterpret_cast<const uint4 *>(src + a0);
        uint4 *__restrict__ d4 = dst ? reinterpret_cast<uint4 *>(dst + a0) : nullptr;
        const unsigned int k_body = (unsigned int)((a0 >> 2) - base_word);

        long long v = tid;
        if constexpr (kWide) {
            // 32-byte vectors when the body starts 32-byte aligned (every whole unit does) and is not a multicast unit
            if (((a0 & 31) == 0) && !mc) {
                const long long n32 = (a1 - a0) >> 5;
                const uint8_t *sb = src + a0;
                uint8_t *db = dst ? dst + a0 : nullptr;
                long long w = tid;
                for (; w + (long long)(kUnroll - 1) * kThreads < n32; w += (long long)kUnroll * kThreads) {
                    U8 r[kUnroll];
#pragma unroll
                    for (int j = 0; j < kUnroll; ++j) r[j] = ld_stream32(sb + ((w + j * kThreads) << 5));
                    if (db) {
#pragma unroll
                        for (int j = 0; j < kUnroll; ++j) st_stream32(db + ((w + j * kThreads) << 5), r[j]);
                    }
#pragma unroll
                    for (int j = 0; j < kUnroll; ++j) {
                        const unsigned int k0 = k_body + 8u * (unsigned int)(w + j * kThreads);
                        acc.vec(r[j].lo, k0);
                        acc.vec(r[j].hi, k0 + 4u);
                    }
                }
                for (; w < n32; w += kThreads) {
                    const U8 r = ld_stream32(sb + (w << 5));
                    if (db) st_stream32(db + (w << 5), r);
                    const unsigned int k0 = k_body + 8u * (unsigned int)w;
                    acc.vec(r.lo, k0);
                    acc.vec(r.hi, k0 + 4u);
                }
                v = 2 * n32 + tid;                       // at most one 16-byte vector is left for the loop below
            }
        }
        // full tiles: kUnroll independent 16-byte loads in flight per thread, then the stores
        for (; v + (long long)(kUnroll - 1) * kThreads < nvec; v += (long long)kUnroll * kThreads) {
            uint4 r[kUnroll];
            if (mc) {
#pragma unroll
                for (int j = 0; j < kUnroll; ++j) r[j] = ld_plain(s4 + v + j * kThreads);
            } else {
#pragma unroll
                for (int j = 0; j < kUnroll; ++j) r[j] = ld_stream(s4 + v + j * kThreads);
            }
            if (d4) {
                if (mc) {
#pragma unroll
                    for (int j = 0; j < kUnroll; ++j) st_multicast(d4 + v + j * kThreads, r[j]);
                } else {
#pragma unroll
                    for (int j = 0; j < kUnroll; ++j) st_stream(d4 + v + j * kThreads, r[j]);
                }
            }
#pragma unroll
            for (int j = 0; j < kUnroll; ++j) acc.vec(r[j], k_body + 4u * (unsigned int)(v + j * kThreads));
        }
        for (; v < nvec; v += kThreads) {
            const uint4 r = mc ? ld_plain(s4 + v) : ld_stream(s4 + v);
            if (d4) { if (mc) st_multicast(d4 + v, r); else st_stream(d4 + v, r); }
            acc.vec(r, k_body + 4u * (unsigned int)v);
        }

        // block reduction; the barrier also orders every thread's stores before thread 0's release below
        const unsigned long long wa = warp_sum(acc.a), ws = warp_sum(acc.s2);
        if (lane == 0) { s_red[p][0][warp] = wa; s_red[p][1][warp] = ws; }
        __syncthreads();
        if (warp == 0) {
            unsigned long long A = lane < kWarps ? s_red[p][0][lane] : 0ull;
            unsigned long long S2 = lane < kWarps ? s_red[p][1][lane] : 0ull;
            A = warp_sum(A);
            S2 = warp_sum(S2);
            if (lane == 0) {
                const unsigned long long B = (unsigned long long)(base_word + 1) * A + S2;
                unsigned long long *cs = D->csum;
                if (cs) {
                    const long long c = b0 >> kChunkShift;
                    long long ce = (c + 1) << kChunkShift;
                    const long long total = __ldg(&D->total);
                    if (ce > total) ce = total;
                    // the unit is its chunk (kFlagStoreSums): the one and only contribution -> plain stores, no prior zeroing
                    const bool whole = (__ldg(&D->flags) & kFlagStoreSums) && b0 == (c << kChunkShift) && b1 == ce;
                    if (whole) {
                        if (mc) { st_multicast_u64(cs + 2 * c, A); st_multicast_u64(cs + 2 * c + 1, B); }
                        else { cs[2 * c] = A; cs[2 * c + 1] = B; }
                    } else if (mc) { red_multicast_u64(cs + 2 * c, A); red_multicast_u64(cs + 2 * c + 1, B); }
                    else { atomicAdd(cs + 2 * c, A); atomicAdd(cs + 2 * c + 1, B); }
                }
                unsigned long long *rcv = D->received;
                if (rcv) {
                    __threadfence_system();       // data and checksum before the counter, at system scope
                    if (mc) red_multicast_u64(rcv, (unsigned long long)(b1 - b0));
                    else atomicAdd_system(rcv, (unsigned long long)(b1 - b0));
                }
            }
        }
    }

    // re-arm the counters for the next launch: the last CTA to drain resets them
    if (tid == 0) {
        __threadfence();
        const unsigned int prev = atomicAdd(&counters->done_ctas, 1u);
        if (prev == gridDim.x - 1) {
            counters->ticket = 0ull;
            counters->done_ctas = 0u;
            __threadfence();
        }
    }
}

// ------------------------------------------------------------------------------------------------
// zero the checksum entries of fully covered chunks (one warp per descriptor, grid-stride)
// ------------------------------------------------------------------------------------------------
__global__ void zero_csum_kernel(const XferDesc *__restrict__ descs, int n_descs) {
    const int lane = threadIdx.x & 31;
    const int warps_per_cta = blockDim.x >> 5;
    for (int d = blockIdx.x * warps_per_cta + (threadIdx.x >> 5); d < n_descs; d += gridDim.x * warps_per_cta) {
        const XferDesc *D = descs + d;
        unsigned long long *cs = D->csum;
        if (!cs) continue;
        const long long off = D->off, end = D->off + D->size, total = D->total;
        const bool mc = (D->flags & kFlagMulticast) != 0;
        const bool stored = (D->flags & kFlagStoreSums) != 0;                   // whole chunks are stored, not accumulated
        const long long c0 = (off + kChunkBytes - 1) >> kChunkShift;            // first chunk starting inside
        for (long long c = c0 + lane; !stored && (c << kChunkShift) < end; c += 32) {
            long long ce = (c + 1) << kChunkShift;
            if (ce > total) ce = total;
            if (ce <= end) {
                if (mc) { st_multicast_u64(cs + 2 * c, 0ull); st_multicast_u64(cs + 2 * c + 1, 0ull); }
                else { cs[2 * c] = 0ull; cs[2 * c + 1] = 0ull; }
            }
        }
        if (lane == 0) {
            // partially covered chunks this descriptor owns
            const long long ch = off >> kChunkShift, ct = (end - 1) >> kChunkShift;
            long long che = (ch + 1) << kChunkShift, cte = (ct + 1) << kChunkShift;
            if (che > total) che = total;
            if (cte > total) cte = total;
            const bool head_partial = (off & (kChunkBytes - 1)) != 0 || end < che;
            const bool tail_partial = end < cte || ((ct << kChunkShift) < off);
            if ((D->flags & kFlagZeroHead) && head_partial) {
                if (mc) { st_multicast_u64(cs + 2 * ch, 0ull); st_multicast_u64(cs + 2 * ch + 1, 0ull); }
                else { cs[2 * ch] = 0ull; cs[2 * ch + 1] = 0ull; }
            }
            if ((D->flags & kFlagZeroTail) && tail_partial && ct != ch) {
                if (mc) { st_multicast_u64(cs + 2 * ct, 0ull); st_multicast_u64(cs + 2 * ct + 1, 0ull); }
                else { cs[2 * ct] = 0ull; cs[2 * ct + 1] = 0ull; }
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// checksum tables of several layers -> one contiguous buffer (one CTA per table, grid-stride)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
gather_tables_kernel(const GatherItem *__restrict__ items, int n_items, unsigned long long *__restrict__ dst) {
    for (int it = blockIdx.x; it < n_items; it += gridDim.x) {
        const unsigned long long *src = items[it].src;
        unsigned long long *out = dst + items[it].dst_word;
        const long long words = items[it].words;
        for (long long w = threadIdx.x; w < words; w += blockDim.x) out[w] = src[w];
    }
}

cudaError_t launch_gather_tables(const GatherItem *items, int n_items, unsigned long long *dst, int n_sm, cudaStream_t stream) {
    if (n_items <= 0) return cudaSuccess;
    const int grid = n_items < 4 * n_sm ? n_items : 4 * n_sm;
    gather_tables_kernel<<<grid, 256, 0, stream>>>(items, n_items, dst);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// K6: synthetic layer contents
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long mix64(unsigned long long z) {
    z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull;
    z ^= z >> 27; z *= 0x94D049BB133111EBull;
    z ^= z >> 31;
    return z;
}

__global__ void __launch_bounds__(256)
fill_pattern_kernel(uint8_t *__restrict__ dst, long long size, unsigned long long base) {
    const long long n2 = size >> 4;                              // pairs of u64 words
    const long long stride = (long long)gridDim.x * blockDim.x;
    ulonglong2 *d2 = reinterpret_cast<ulonglong2 *>(dst);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += stride) {
        ulonglong2 w;
        w.x = mix64(base + 2ull * i);
        w.y = mix64(base + 2ull * i + 1ull);
        d2[i] = w;
    }
    if (blockIdx.x == 0 && threadIdx.x < (size & 15)) {          // trailing bytes
        const long long pos = (n2 << 4) + threadIdx.x;
        const unsigned long long w = mix64(base + (unsigned long long)(pos >> 3));
        dst[pos] = (uint8_t)(w >> (8 * (pos & 7)));
    }
}

// ------------------------------------------------------------------------------------------------
// launch wrappers
// ------------------------------------------------------------------------------------------------
cudaError_t launch_zero_csum(const XferDesc *descs, int n_descs, cudaStream_t stream) {
    if (n_descs <= 0) return cudaSuccess;
    const int warps_per_cta = 8;
    int grid = (n_descs + warps_per_cta - 1) / warps_per_cta;
    if (grid > 1024) grid = 1024;
    zero_csum_kernel<<<grid, warps_per_cta * 32, 0, stream>>>(descs, n_descs);
    return cudaGetLastError();
}

template <int kThreads, int kUnroll, bool kWide = false>
static cudaError_t launch_reg(const XferDesc *descs, int n_descs, long long total_units, int unit_shift,
                              SchedCounters *counters, const LaunchCfg &cfg, cudaStream_t stream,
                              int *grid_out) {
    int per_sm = cfg.ctas_per_sm;
    if (per_sm <= 0) {
        cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(
            &per_sm, copy_csum_kernel<kThreads, kUnroll, kWide>, kThreads, 0);
        if (e != cudaSuccess) return e;
        if (per_sm < 1) per_sm = 1;
    }
    long long grid = (long long)cfg.n_sm * per_sm;
    if (grid > total_units) grid = total_units;
    if (grid < 1) grid = 1;
    if (grid_out) *grid_out = (int)grid;
    copy_csum_kernel<kThreads, kUnroll, kWide><<<(int)grid, kThreads, 0, stream>>>(descs, n_descs, total_units,
                                                                            unit_shift, counters);
    return cudaGetLastError();
}

cudaError_t launch_copy_csum_tma(const XferDesc *descs, int n_descs, long long total_units, int unit_shift,
                                 SchedCounters *counters, const LaunchCfg &cfg, cudaStream_t stream,
                                 int *grid_out);   // dissem_kernels_tma.cu

cudaError_t launch_copy_csum(const XferDesc *descs, int n_descs, long long total_units, int unit_shift,
                             SchedCounters *counters, const LaunchCfg &cfg, cudaStream_t stream,
                             int *grid_out) {
    if (total_units <= 0) { if (grid_out) *grid_out = 0; return cudaSuccess; }
    if (cfg.variant == 1)
        return launch_copy_csum_tma(descs, n_descs, total_units, unit_shift, counters, cfg, stream, grid_out);
    switch (cfg.reg_cfg) {
        case 0: return launch_reg<256, 4>(descs, n_descs, total_units, unit_shift, counters, cfg, stream, grid_out);
        case 2: return launch_reg<256, 8>(descs, n_descs, total_units, unit_shift, counters, cfg, stream, grid_out);
        case 3: return launch_reg<512, 8>(descs, n_descs, total_units, unit_shift, counters, cfg, stream, grid_out);
        case 4: return launch_reg<1024, 4>(descs, n_descs, total_units, unit_shift, counters, cfg, stream, grid_out);
        case 5: return launch_reg<1024, 2>(descs, n_descs, total_units, unit_shift, counters, cfg, stream, grid_out);
        case 6: return launch_reg<512, 2>(descs, n_descs, total_units, unit_shift, counters, cfg, stream, grid_out);
        case 7: return launch_reg<512, 2, true>(descs, n_descs, total_units, unit_shift, counters, cfg, stream, grid_out);
        case 8: return launch_reg<512, 4, true>(descs, n_descs, total_units, unit_shift, counters, cfg, stream, grid_out);
        case 9: return launch_reg<256, 4, true>(descs, n_descs, total_units, unit_shift, counters, cfg, stream, grid_out);
        default: return launch_reg<512, 4>(descs, n_descs, total_units, unit_shift, counters, cfg, stream, grid_out);
    }
}

cudaError_t warm_kernels_tma();   // dissem_kernels_tma.cu

cudaError_t warm_kernels() {
    cudaFuncAttributes a;
    cudaError_t e = cudaFuncGetAttributes(&a, copy_csum_kernel<512, 4>);
    if (e == cudaSuccess) e = cudaFuncGetAttributes(&a, zero_csum_kernel);
    if (e == cudaSuccess) e = cudaFuncGetAttributes(&a, fill_pattern_kernel);
    if (e == cudaSuccess) e = cudaFuncGetAttributes(&a, gather_tables_kernel);
    if (e == cudaSuccess) e = warm_kernels_tma();
    return e;
}

cudaError_t launch_fill_pattern(uint8_t *dst, long long size, unsigned long long layer_id,
                                unsigned long long seed, int n_sm, cudaStream_t stream) {
    if (size <= 0) return cudaSuccess;
    const unsigned long long base = seed + (layer_id + 1ull) * 0x9E3779B97F4A7C15ull;
    long long want = ((size >> 4) + 255) / 256;
    long long cap = (long long)n_sm * 8;
    int grid = (int)(want < 1 ? 1 : (want > cap ? cap : want));
    fill_pattern_kernel<<<grid, 256, 0, stream>>>(dst, size, base);
    return cudaGetLastError();
}

}  // namespace dissem
