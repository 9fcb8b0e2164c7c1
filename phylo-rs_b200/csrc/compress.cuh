// compress.cuh — site-pattern compression on the device: Alignment::compress_columns
// (src/alignment.rs:93-125) producing CompressedColumns {patterns, site_to_pattern, multiplicity}
// (:129-138) with the caller's row order as leaf_order.
//
// The reference walks the sites in order and looks every column (a Vec<u8> over ALL rows, raw bytes —
// 'T' and 'U' are different keys, rows that name internal nodes take part) up in a HashMap; a new
// column gets the next pattern index, so patterns are numbered in order of first appearance.
// Here:
//   column_hash_kernel      64-bit keyed hash of every column, accumulated row chunk by row chunk
//                           while the next chunk is still in flight from the host (HBM-bound: reads
//                           rows x width bytes once, 8 sites per thread, 8-byte coalesced loads)
//   stable radix sort of (hash, site) + head flags + scans (CUB for the sort/scan primitives)
//   column_verify_kernel    every site that is not the first of its hash group is compared byte for
//                           byte with that first site, so equal hashes never merge different columns:
//                           a mismatch re-runs the whole thing under another hash key
//   pattern ids             exclusive scan of the "first of its group" flags in SITE order
//                           = the reference's first-appearance numbering, exactly
//   gather_patterns_kernel  unique columns -> tip codes [row][pattern] (the layout every kernel of
//                           the path reads), per-row invalid-byte flags, multiplicities
#pragma once
#include "kernels.cuh"

#ifndef PHB_HOST_EMU
#include <cub/cub.cuh>
#else
#include <algorithm>
#include <numeric>
#include <vector>
#endif

namespace phb {

constexpr int HASH_SITES_PER_THREAD = 8;

__host__ __device__ __forceinline__ uint64_t hash_step(uint64_t h, uint32_t byte, uint64_t mul)
{
    h = ((h << 5) | (h >> 59)) ^ (uint64_t)byte;
    return h * mul;
}
__host__ __device__ __forceinline__ uint64_t hash_final(uint64_t h)
{
    h ^= h >> 33;
    h *= 0xFF51AFD7ED558CCDull;
    h ^= h >> 33;
    h *= 0xC4CEB9FE1A85EC53ull;
    return h ^ (h >> 33);
}

// raw: [n_rows][pitch] bytes (pitch a multiple of 16); rows [r0, r0 + nr) are folded into hash[site].
// first != 0 starts from the key; last != 0 applies the finaliser.
__global__ void __launch_bounds__(256)
column_hash_kernel(const uint8_t *__restrict__ raw, int64_t pitch, int r0, int nr, int64_t width, uint64_t key,
                   uint64_t mul, int first, int last, uint64_t *__restrict__ hash)
{
    const int64_t s0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * HASH_SITES_PER_THREAD;
    if (s0 >= width) return;
    uint64_t h[HASH_SITES_PER_THREAD];
#pragma unroll
    for (int k = 0; k < HASH_SITES_PER_THREAD; ++k) h[k] = first ? key : (s0 + k < width ? hash[s0 + k] : 0);
    const uint8_t *p = raw + (int64_t)r0 * pitch + s0;
    for (int r = 0; r < nr; ++r, p += pitch) {
        const uint2 w = *reinterpret_cast<const uint2 *>(p);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            h[k] = hash_step(h[k], (w.x >> (8 * k)) & 0xFFu, mul);
            h[4 + k] = hash_step(h[4 + k], (w.y >> (8 * k)) & 0xFFu, mul);
        }
    }
#pragma unroll
    for (int k = 0; k < HASH_SITES_PER_THREAD; ++k)
        if (s0 + k < width) hash[s0 + k] = last ? hash_final(h[k]) : h[k];
}

__global__ void iota_kernel(uint32_t *__restrict__ v, int64_t n)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) v[i] = (uint32_t)i;
}

// sorted position i starts a hash group when its key differs from its predecessor's
__global__ void group_head_kernel(const uint64_t *__restrict__ keys, int64_t n, uint32_t *__restrict__ head)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) head[i] = (i == 0 || keys[i] != keys[i - 1]) ? 1u : 0u;
}

// gid: inclusive scan of head (1-based group id per sorted position) -> group_start[g]
__global__ void group_start_kernel(const uint32_t *__restrict__ head, const uint32_t *__restrict__ gid, int64_t n,
                                   uint32_t *__restrict__ group_start)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && head[i]) group_start[gid[i] - 1] = (uint32_t)i;
}

// rep[site] = first (smallest, the sort is stable) site of the site's group; is_rep[site] = (rep == site);
// group_size_at_rep[rep] = number of sites in the group
__global__ void group_rep_kernel(const uint32_t *__restrict__ sorted_sites, const uint32_t *__restrict__ gid,
                                 const uint32_t *__restrict__ group_start, uint32_t n_groups, int64_t n,
                                 uint32_t *__restrict__ rep, uint32_t *__restrict__ is_rep,
                                 uint32_t *__restrict__ size_at_rep)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t g = gid[i] - 1;
    const uint32_t st = group_start[g];
    const uint32_t en = (g + 1 < n_groups) ? group_start[g + 1] : (uint32_t)n;
    const uint32_t s = sorted_sites[i], r = sorted_sites[st];
    rep[s] = r;
    is_rep[s] = (s == r) ? 1u : 0u;
    if (s == r) size_at_rep[s] = en - st;
}

// exact comparison of every non-first site with the first site of its group over rows [r0, r0 + nr)
__global__ void __launch_bounds__(256)
column_verify_kernel(const uint8_t *__restrict__ raw, int64_t pitch, int n_rows, int rows_per_block, int64_t width,
                     const uint32_t *__restrict__ rep, unsigned int *__restrict__ mismatch)
{
    const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= width) return;
    const uint32_t r = rep[s];
    if (r == (uint32_t)s) return;
    const int r0 = blockIdx.y * rows_per_block;
    const int r1 = min(n_rows, r0 + rows_per_block);
    bool bad = false;
    for (int row = r0; row < r1; ++row) bad |= raw[(int64_t)row * pitch + s] != raw[(int64_t)row * pitch + r];
    if (bad) atomicOr(mismatch, 1u);
}

// pid: exclusive scan of is_rep in site order.  rep_list[pattern] = site, s2p[site] = pattern, mult[pattern].
__global__ void pattern_index_kernel(const uint32_t *__restrict__ rep, const uint32_t *__restrict__ is_rep,
                                     const uint32_t *__restrict__ pid, const uint32_t *__restrict__ size_at_rep,
                                     int64_t width, uint32_t *__restrict__ rep_list, uint64_t *__restrict__ s2p,
                                     uint64_t *__restrict__ mult)
{
    const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= width) return;
    const uint32_t r = rep[s];
    s2p[s] = pid[r];
    if (is_rep[s]) {
        rep_list[pid[s]] = (uint32_t)s;
        mult[pid[s]] = size_at_rep[s];
    }
}

// unique columns [p0, p0 + np) -> tip codes of one device shard: tips[row][dst_off + (p - p0)].
// One thread = 4 consecutive patterns x GATHER_ROWS rows: the four column indices are read once and stay in
// registers, every row costs four byte loads (neighbouring threads read neighbouring columns whenever the
// unique columns are mostly in site order) and one 32-bit store.  P_pad and dst_off are multiples of 4.
constexpr int GATHER_ROWS = 16;
__global__ void __launch_bounds__(256)
gather_patterns_kernel(const uint8_t *__restrict__ raw, int64_t pitch, const uint32_t *__restrict__ rep_list, int64_t p0,
                       int64_t np, int n_rows, uint8_t *__restrict__ tips, int64_t P_pad, int64_t dst_off,
                       int *__restrict__ row_invalid)
{
    __shared__ uint8_t lut[256]; // byte -> tip code (blockDim.x == 256)
    lut[threadIdx.x] = encode_byte((uint8_t)threadIdx.x);
    __syncthreads();
    const int64_t p = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (p >= np) return;
    const int n = (int)min((int64_t)4, np - p);
    uint32_t col[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) col[k] = rep_list[p0 + p + (k < n ? k : 0)];
    const int r0 = blockIdx.y * GATHER_ROWS, r1 = min(n_rows, r0 + GATHER_ROWS);
    for (int r = r0; r < r1; ++r) {
        const uint8_t *src = raw + (int64_t)r * pitch;
        uint32_t c[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) c[k] = lut[src[col[k]]];
        uint8_t *dst = tips + (int64_t)r * P_pad + dst_off + p;
        bool bad = false;
        if (n == 4) {
            *reinterpret_cast<uint32_t *>(dst) = c[0] | (c[1] << 8) | (c[2] << 16) | (c[3] << 24);
            bad = c[0] == CODE_INVALID || c[1] == CODE_INVALID || c[2] == CODE_INVALID || c[3] == CODE_INVALID;
        } else {
            for (int k = 0; k < n; ++k) {
                dst[k] = (uint8_t)c[k];
                bad |= c[k] == CODE_INVALID;
            }
        }
        if (bad) row_invalid[r] = 1;
    }
}

__global__ void mult_u64_to_double_kernel(const uint64_t *__restrict__ in, int64_t p0, int64_t n, int64_t n_pad,
                                          double *__restrict__ out)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_pad; i += (int64_t)gridDim.x * blockDim.x)
        out[i] = (i < n) ? (double)in[p0 + i] : 0.0;
}

// ---- sort / scan primitives -----------------------------------------------------------------------
// Product build: CUB device-wide radix sort (stable) and prefix sums on the context's stream.
// Emulation build (tests/emu, "device" memory is host memory): the std:: equivalents.
inline size_t compress_temp_bytes(int64_t n)
{
#ifndef PHB_HOST_EMU
    size_t a = 0, b = 0, c = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, a, (const uint64_t *)nullptr, (uint64_t *)nullptr, (const uint32_t *)nullptr,
                                    (uint32_t *)nullptr, (int)n);
    cub::DeviceScan::InclusiveSum(nullptr, b, (const uint32_t *)nullptr, (uint32_t *)nullptr, (int)n);
    cub::DeviceScan::ExclusiveSum(nullptr, c, (const uint32_t *)nullptr, (uint32_t *)nullptr, (int)n);
    return std::max(a, std::max(b, c));
#else
    (void)n;
    return 16;
#endif
}

inline cudaError_t dev_sort_pairs(void *temp, size_t temp_bytes, const uint64_t *keys_in, uint64_t *keys_out,
                                  const uint32_t *vals_in, uint32_t *vals_out, int64_t n, cudaStream_t stream)
{
#ifndef PHB_HOST_EMU
    return cub::DeviceRadixSort::SortPairs(temp, temp_bytes, keys_in, keys_out, vals_in, vals_out, (int)n, 0, 64, stream);
#else
    (void)temp; (void)temp_bytes; (void)stream;
    std::vector<uint32_t> idx((size_t)n);
    std::iota(idx.begin(), idx.end(), 0u);
    std::stable_sort(idx.begin(), idx.end(), [&](uint32_t a, uint32_t b) { return keys_in[a] < keys_in[b]; });
    for (int64_t i = 0; i < n; ++i) {
        keys_out[i] = keys_in[idx[(size_t)i]];
        vals_out[i] = vals_in[idx[(size_t)i]];
    }
    return cudaSuccess;
#endif
}

inline cudaError_t dev_scan(void *temp, size_t temp_bytes, const uint32_t *in, uint32_t *out, int64_t n, bool inclusive,
                            cudaStream_t stream)
{
#ifndef PHB_HOST_EMU
    return inclusive ? cub::DeviceScan::InclusiveSum(temp, temp_bytes, in, out, (int)n, stream)
                     : cub::DeviceScan::ExclusiveSum(temp, temp_bytes, in, out, (int)n, stream);
#else
    (void)temp; (void)temp_bytes; (void)stream;
    uint32_t acc = 0;
    for (int64_t i = 0; i < n; ++i) {
        const uint32_t v = in[i];
        out[i] = inclusive ? acc + v : acc;
        acc += v;
    }
    return cudaSuccess;
#endif
}

} // namespace phb
