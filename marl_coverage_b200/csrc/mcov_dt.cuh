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

// A team = a contiguous range of warps of the CTA that runs the transform together (the whole CTA, or
// a subset while other warps stream output).  Named barrier `bar` (0 = __syncthreads' barrier; use it
// only when the team is the whole CTA).
struct Team {
    int rank, size, bar;
    __device__ __forceinline__ void sync() const {
        asm volatile("bar.sync %0, %1;" :: "r"(bar), "r"(size) : "memory");
    }
    __device__ __forceinline__ int any(int pred) const {
        int r;
        asm volatile("{\n\t.reg .pred p, q;\n\tsetp.ne.s32 p, %3, 0;\n\t"
                     "bar.red.or.pred q, %1, %2, p;\n\tselp.s32 %0, 1, 0, q;\n\t}"
                     : "=r"(r) : "r"(bar), "r"(size), "r"(pred) : "memory");
        return r;
    }
};
__device__ __forceinline__ Team whole_cta() { Team t; t.rank = threadIdx.x; t.size = blockDim.x; t.bar = 0; return t; }

// Shared scratch the transform needs besides the distance buffer (ints).
__host__ __device__ inline int dt_scratch_ints(int L, int threads) {
    const int lp = (L + 1) / 2;
    return 2 + 2 * (lp > threads ? lp : threads);
}

// d[x * pitch + y] = taxicab distance from (x, y) to the nearest ZERO pixel (0 on zero pixels).
// `zero_word(strip, x)` returns the zero-pixel bits of one plane word (bit b <-> y = 32*strip + b);
// bits with y >= L must be 0.
//   pass 1 (along y): closed form on the bit words with clz / ffs — no sequential scan;
//   pass 2 (along x): min-plus sweep on PAIRS of columns (two uint16 per 32-bit lane, vminu2 /
//     saturating vaddus2), each column pair cut into `nch` chunks so that the whole team works:
//     local forward+backward scan per chunk, chunk-end values combined across chunks (min-plus is
//     associative: entering a chunk with carry c adds c + offset), then applied in a second sweep.
// Every thread of the team must call; `red` is dt_scratch_ints(L, team.size) ints of shared scratch.
template <class ZeroWord>
__device__ DtResult l1_transform(ZeroWord zero_word, const Geometry& g, uint16_t* d, int pitch, int* red,
                                 const Team team) {
    constexpr unsigned FULLM = 0xffffffffu;
    constexpr unsigned INF2 = 0xffffffffu;                   // two DT_INF halves
    const int W = g.W, L = g.L, strips = g.strips;
    const int T = team.size, tid = team.rank;
    if (tid == 0) { red[0] = 0; red[1] = 0; }
    int anyz = 0;
    const int nw = W * strips;
    for (int idx = tid; idx < nw; idx += T) {
        const int s = idx / W, x = idx - s * W;              // x fastest within a strip
        const uint32_t z = zero_word(s, x);
        anyz |= (z != 0u);
        uint32_t* drow = reinterpret_cast<uint32_t*>(d + (size_t)x * pitch + 32 * s);
        if (z == FULLM) {                                    // all zero pixels (obstacles / not yet covered)
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
    anyz = team.any(anyz);

    // ---- pass 2 ----------------------------------------------------------------------------------
    const int LP = (L + 1) >> 1;                             // column pairs (pitch is even and > L)
    const int ph = pitch >> 1;                               // 32-bit words per x row
    int nch = T / LP;
    if (nch < 1) nch = 1;
    if (nch > W) nch = W;
    const int K = (W + nch - 1) / nch;                       // rows per chunk
    uint32_t* endF = reinterpret_cast<uint32_t*>(red + 2);   // [nch][LP] forward value at the chunk's last row
    uint32_t* endB = endF + (LP > T ? LP : T);               // [nch][LP] backward value at the chunk's first row
    const int items = LP * nch;
    constexpr unsigned ONE2 = 0x00010001u;
    uint32_t* d32 = reinterpret_cast<uint32_t*>(d);
    for (int it = tid; it < items; it += T) {
        const int c = it / LP, yp = it - c * LP;
        const int x0 = c * K, x1 = min(x0 + K, W);
        uint32_t* col = d32 + yp;
        uint32_t v = INF2;
        int x = x0;
        for (; x + 4 <= x1; x += 4) {
            uint32_t t0 = col[(size_t)x * ph], t1 = col[(size_t)(x + 1) * ph];
            uint32_t t2 = col[(size_t)(x + 2) * ph], t3 = col[(size_t)(x + 3) * ph];
            v = __vminu2(t0, __vaddus2(v, ONE2)); col[(size_t)x * ph] = v;
            v = __vminu2(t1, __vaddus2(v, ONE2)); col[(size_t)(x + 1) * ph] = v;
            v = __vminu2(t2, __vaddus2(v, ONE2)); col[(size_t)(x + 2) * ph] = v;
            v = __vminu2(t3, __vaddus2(v, ONE2)); col[(size_t)(x + 3) * ph] = v;
        }
        for (; x < x1; ++x) { v = __vminu2(col[(size_t)x * ph], __vaddus2(v, ONE2)); col[(size_t)x * ph] = v; }
        endF[it] = v;                                        // INF2 for an empty chunk
        v = INF2;
        x = x1 - 1;
        for (; x - 3 >= x0; x -= 4) {
            uint32_t t0 = col[(size_t)x * ph], t1 = col[(size_t)(x - 1) * ph];
            uint32_t t2 = col[(size_t)(x - 2) * ph], t3 = col[(size_t)(x - 3) * ph];
            v = __vminu2(t0, __vaddus2(v, ONE2)); col[(size_t)x * ph] = v;
            v = __vminu2(t1, __vaddus2(v, ONE2)); col[(size_t)(x - 1) * ph] = v;
            v = __vminu2(t2, __vaddus2(v, ONE2)); col[(size_t)(x - 2) * ph] = v;
            v = __vminu2(t3, __vaddus2(v, ONE2)); col[(size_t)(x - 3) * ph] = v;
        }
        for (; x >= x0; --x) { v = __vminu2(col[(size_t)x * ph], __vaddus2(v, ONE2)); col[(size_t)x * ph] = v; }
        endB[it] = v;
    }
    team.sync();
    uint32_t mx2 = 0;
    for (int it = tid; it < items; it += T) {
        const int c = it / LP, yp = it - c * LP;
        const int x0 = c * K, x1 = min(x0 + K, W);
        // carry entering from the left = global forward value at row x0 - 1; from the right likewise
        uint32_t cf = INF2, cb = INF2;
        for (int cc = 0; cc < c; ++cc) {
            const uint32_t len = (uint32_t)(min(cc * K + K, W) - cc * K);
            cf = __vminu2(endF[cc * LP + yp], __vaddus2(cf, len | (len << 16)));
        }
        for (int cc = nch - 1; cc > c; --cc) {
            const int a0 = cc * K, a1 = min(a0 + K, W);
            if (a1 > a0) {
                const uint32_t len = (uint32_t)(a1 - a0);
                cb = __vminu2(endB[cc * LP + yp], __vaddus2(cb, len | (len << 16)));
            }
        }
        uint32_t* col = d32 + yp;
        const uint32_t keep = (2 * yp + 1 < L) ? 0xffffffffu : 0x0000ffffu;   // odd L: last pair's high half is padding
        uint32_t up = __vaddus2(cf, ONE2);                                    // cf + (x - x0 + 1)
        for (int x = x0; x < x1; ++x) {
            const uint32_t k = (uint32_t)(x1 - x);
            const uint32_t down = __vaddus2(cb, k | (k << 16));               // cb + (x1 - x)
            uint32_t v = col[(size_t)x * ph];
            v = __vminu2(v, __vminu2(up, down));
            col[(size_t)x * ph] = v;
            mx2 = __vmaxu2(mx2, v & keep);
            up = __vaddus2(up, ONE2);
        }
    }
    int mx = (int)max(mx2 & 0xffffu, mx2 >> 16);
    mx = __reduce_max_sync(FULLM, mx);
    if ((tid & 31) == 0 && mx > 0) atomicMax(&red[1], mx);
    team.sync();
    DtResult r;
    r.any_zero = anyz;
    r.dmax = red[1];
    team.sync();
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
