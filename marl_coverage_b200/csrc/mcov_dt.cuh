// mcov_dt.cuh — exact L1 (taxicab) distance transform on strip-major bit planes, CTA-wide.
// Restates cv2.distanceTransform(img, cv2.DIST_L1, cv2.DIST_MASK_PRECISE) as the reference calls it
// (super_grid_rl.py:295, dec_grid_rl.py:274): OpenCV is an un-vendored dependency of the reference;
// the L1 transform is exact integer arithmetic, hence version independent.
#pragma once
#include "mcov_common.cuh"

namespace mcov {

constexpr int DT_INF = 0xffff;                // "no zero pixel in reach" (> W + L for any int16 map)

struct DtResult {
    int any_zero;                             // the image has at least one zero pixel
    int dmax;                                 // max distance over the image (valid when any_zero)
};

// uint16 elements per x row of the work buffer: 2 * (odd number of 32-bit words), so that the pass
// along y (threads = x), the pass along x (threads = y) and row readers all hit distinct banks.
__host__ __device__ inline int dt_pitch_of(const Geometry& g) {
    int words = g.strips * 16;
    if ((words & 1) == 0) words += 1;
    return 2 * words;
}

// Shared scratch the transform needs besides the distance buffer: 2 ints + 2 * max(blockDim, L) ints.
__host__ __device__ inline int dt_scratch_ints(int L, int threads) { return 2 + 2 * (L > threads ? L : threads); }

// d[x * pitch + y] = taxicab distance from (x, y) to the nearest ZERO pixel (0 on zero pixels).
// `zero_word(strip, x)` returns the zero-pixel bits of one plane word (bit b <-> y = 32*strip + b);
// bits with y >= L must be 0.
//   pass 1 (along y): closed form on the bit words with clz / ffs — no sequential scan;
//   pass 2 (along x): min-plus sweep, each column cut into `nch` chunks so that L * nch threads work:
//     local forward+backward scan per chunk, chunk-end values combined across chunks (min-plus is
//     associative: entering a chunk with carry c adds c + offset), then applied in a second sweep.
// All threads of the CTA must call; `red` is dt_scratch_ints(L, blockDim.x) ints of shared scratch.
template <class ZeroWord>
__device__ DtResult l1_transform(ZeroWord zero_word, const Geometry& g, uint16_t* d, int pitch, int* red) {
    constexpr unsigned FULLM = 0xffffffffu;
    const int W = g.W, L = g.L, strips = g.strips;
    const int T = blockDim.x;
    if (threadIdx.x == 0) { red[0] = 0; red[1] = 0; }
    __syncthreads();
    int anyz = 0;
    const int nw = W * strips;
    for (int idx = threadIdx.x; idx < nw; idx += T) {
        const int s = idx / W, x = idx - s * W;              // x fastest within a strip
        const uint32_t z = zero_word(s, x);
        anyz |= (z != 0u);
        uint32_t* drow = reinterpret_cast<uint32_t*>(d + (size_t)x * pitch + 32 * s);
        if (z == FULLM) {                                    // common early in an episode: all zero pixels
#pragma unroll
            for (int j = 0; j < 16; ++j) drow[j] = 0u;
            continue;
        }
        int base_lo = DT_INF, base_hi = DT_INF;              // gap to the nearest zero pixel in other strips
        for (int ss = s - 1; ss >= 0; --ss) {
            const uint32_t t = zero_word(ss, x);
            if (t) { base_lo = 32 * (s - ss) - (31 - __clz(t)); break; }
        }
        for (int ss = s + 1; ss < strips; ++ss) {
            const uint32_t t = zero_word(ss, x);
            if (t) { base_hi = 32 * (ss - s) + (__ffs(t) - 1) - 31; break; }
        }
#pragma unroll 4
        for (int j = 0; j < 16; ++j) {
            uint32_t packed = 0;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int b = 2 * j + h;
                const uint32_t mlo = z & (FULLM >> (31 - b)), mhi = z & (FULLM << b);
                const int dl = mlo ? b - (31 - __clz(mlo)) : b + base_lo;
                const int dh = mhi ? (__ffs(mhi) - 1) - b : (31 - b) + base_hi;
                const int v = min(min(dl, dh), DT_INF);
                packed |= (uint32_t)v << (16 * h);
            }
            drow[j] = packed;
        }
    }
    anyz = __syncthreads_or(anyz);

    // ---- pass 2 ----------------------------------------------------------------------------------
    int nch = T / L;
    if (nch < 1) nch = 1;
    if (nch > W) nch = W;
    const int K = (W + nch - 1) / nch;                       // rows per chunk
    int* endF = red + 2;                                     // [nch][L] forward value at the chunk's last row
    int* endB = endF + (L > T ? L : T);                      // [nch][L] backward value at the chunk's first row
    const int items = L * nch;
    for (int it = threadIdx.x; it < items; it += T) {
        const int c = it / L, y = it - c * L;
        const int x0 = c * K, x1 = min(x0 + K, W);
        uint16_t* col = d + y;
        int v = DT_INF;
        for (int xb = x0; xb < x1; xb += 8) {
            int t[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) t[k] = (xb + k < x1) ? (int)col[(size_t)(xb + k) * pitch] : DT_INF;
#pragma unroll
            for (int k = 0; k < 8; ++k)
                if (xb + k < x1) {
                    v = min(min(t[k], v + 1), DT_INF);
                    col[(size_t)(xb + k) * pitch] = (uint16_t)v;
                }
        }
        endF[it] = x1 > x0 ? v : DT_INF;
        v = DT_INF;
        for (int xb = x1 - 1; xb >= x0; xb -= 8) {
            int t[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) t[k] = (xb - k >= x0) ? (int)col[(size_t)(xb - k) * pitch] : DT_INF;
#pragma unroll
            for (int k = 0; k < 8; ++k)
                if (xb - k >= x0) {
                    v = min(min(t[k], v + 1), DT_INF);
                    col[(size_t)(xb - k) * pitch] = (uint16_t)v;
                }
        }
        endB[it] = x1 > x0 ? v : DT_INF;
    }
    __syncthreads();
    int mx = 0;
    for (int it = threadIdx.x; it < items; it += T) {
        const int c = it / L, y = it - c * L;
        const int x0 = c * K, x1 = min(x0 + K, W);
        // carry entering from the left = global forward value at row x0 - 1; from the right likewise
        int cf = DT_INF, cb = DT_INF;
        for (int cc = 0; cc < c; ++cc) cf = min(endF[cc * L + y], cf + (min(cc * K + K, W) - cc * K));
        for (int cc = nch - 1; cc > c; --cc) {
            const int a0 = cc * K, a1 = min(a0 + K, W);
            if (a1 > a0) cb = min(endB[cc * L + y], cb + (a1 - a0));
        }
        cf = min(cf, DT_INF);
        cb = min(cb, DT_INF);
        uint16_t* col = d + y;
        for (int x = x0; x < x1; ++x) {
            int v = (int)col[(size_t)x * pitch];
            v = min(v, min(cf + (x - x0 + 1), cb + (x1 - x)));
            v = min(v, DT_INF);
            col[(size_t)x * pitch] = (uint16_t)v;
            mx = max(mx, v);
        }
    }
    mx = __reduce_max_sync(FULLM, mx);
    if ((threadIdx.x & 31) == 0 && mx > 0) atomicMax(&red[1], mx);
    __syncthreads();
    DtResult r;
    r.any_zero = anyz;
    r.dmax = red[1];
    __syncthreads();
    return r;
}

// float32 value of the inverted distance channel: 1 - d / max(d), all in float32
// (super_grid_rl.py:297-303, dec_grid_rl.py:276-282).
__device__ __forceinline__ float dist_value(int dv, int any_zero, int dmax) {
    if (!any_zero) return 0.0f;                               // cv2: FLT_MAX everywhere -> 1 - 1
    if (dmax <= 0) return 1.0f;                               // every pixel is a zero pixel: no division
    return __fsub_rn(1.0f, __fdiv_rn((float)dv, (float)dmax));
}

}  // namespace mcov
