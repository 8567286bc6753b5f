// k_window_features_v3 — compact-code variant: 8 lanes per window (4 windows per warp), one pass of
// the loop body per axis (not unrolled across axes, so the kernel stays inside the instruction cache —
// the v2 kernel lost ~3 of 4 issue slots to stall_no_instruction), samples held in registers, the sum of
// squares for RESULTANT carried across the axis loop, 14 outputs per axis written lane-parallel.
// Discrete decisions (bin index, peak selection, threshold ladder) use exactly the oracle's operations.
#pragma once

__device__ __noinline__ double precise_sqrt(double s) { return sqrt(s); }

template <int KITER, bool EXACT>       // EXACT: L == 8 * KITER (no bounds checks)
__global__ void __launch_bounds__(128, 4)
k_window_features_v3(const double* __restrict__ x, const double* __restrict__ y, const double* __restrict__ z,
                     const int64_t* __restrict__ ts, const int64_t* __restrict__ win_start, int64_t nw, int L,
                     double* __restrict__ out, int64_t ld_out) {
    constexpr int G = 8;
    const int lane = threadIdx.x & 31, sub = lane & (G - 1), grp = lane >> 3;
    const int64_t quad0 = ((int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5));
    const int64_t nquads = (int64_t)gridDim.x * (blockDim.x >> 5);
    const double dL = (double)L;
    const double qnan = __longlong_as_double(0x7ff8000000000000LL);
    __shared__ unsigned s_bins[4][4][10];                  // [warp][window of the warp][bin]
    __shared__ double s_r2[KITER][128];                    // per-thread running x^2+y^2+z^2 (frees KITER*2 registers)
    double* r2 = &s_r2[0][threadIdx.x];                    // element k at r2[k * 128]: conflict-free
    unsigned* bins_w = s_bins[threadIdx.x >> 5][grp];
    const unsigned gmask = 0xffu << (grp * G);

    for (int64_t qd = quad0; qd * 4 < nw; qd += nquads) {
        const int64_t w = qd * 4 + grp;
        const bool live = w < nw;
        const int64_t st = win_start[live ? w : nw - 1];
        double* dst = out + (live ? w : nw - 1) * ld_out;

#pragma unroll 1
        for (int a = 0; a < 3; a++) {
            const double* p = (a == 0 ? x : (a == 1 ? y : z)) + st;
            double v[KITER];
#pragma unroll
            for (int k = 0; k < KITER; k++) v[k] = (EXACT || sub + G * k < L) ? __ldg(p + sub + G * k) : 0.0;
            // ---- pass A ----
            double sum = 0.0, lo = INFINITY, hi = -INFINITY;
#pragma unroll
            for (int k = 0; k < KITER; k++) {
                if (EXACT || sub + G * k < L) {
                    sum += v[k]; lo = fmin(lo, v[k]); hi = fmax(hi, v[k]);
                    r2[k * 128] = a == 0 ? v[k] * v[k] : fma(v[k], v[k], r2[k * 128]);
                }
            }
#pragma unroll
            for (int m = 4; m >= 1; m >>= 1) {
                sum += shfl_xor_d(sum, m);
                lo = fmin(lo, shfl_xor_d(lo, m));
                hi = fmax(hi, shfl_xor_d(hi, m));
            }
            const double mean = sum / dL;
            const double scale = hi > lo ? 10.0 / (hi - lo) : 0.0;
            // ---- pass B: deviations, 10-bin counts (shared-memory counters of this window), local maxima ----
            double sq = 0.0, ab = 0.0, pm = -INFINITY;
            unsigned pk = 0;
            bins_w[sub] = 0; if (sub + G < 10) bins_w[sub + G] = 0;
            __syncwarp();
#pragma unroll
            for (int k = 0; k < KITER; k++) {
                const int i = sub + G * k;
                if (EXACT || i < L) {
                    const double c = v[k];
                    const double d = c - mean;
                    sq = fma(d, d, sq);
                    ab += fabs(d);
                    const int bq = __double2int_rd((c - lo) * scale);
                    atomicAdd(&bins_w[bq >= 9 ? 9 : bq], 1u);
                    // local maximum test without branches: both neighbours are always loaded (clamped at the
                    // window edges, where the result is masked off)
                    const bool interior = i >= 1 && i + 1 < L;
                    const double lft = __ldg(p + (i >= 1 ? i - 1 : i)), rgt = __ldg(p + (i + 1 < L ? i + 1 : i));
                    const bool is_pk = interior & (c > lft) & (c >= rgt);
                    pk |= (is_pk ? 1u : 0u) << k;
                    pm = is_pk ? fmax(pm, c) : pm;
                }
            }
            int npk = __popc(pk);
#pragma unroll
            for (int m = 4; m >= 1; m >>= 1) {
                sq += shfl_xor_d(sq, m);
                ab += shfl_xor_d(ab, m);
                npk += __shfl_xor_sync(0xffffffffu, npk, m);
                pm = fmax(pm, shfl_xor_d(pm, m));
            }
            // ---- peaks: first rung of the ladder tau_k = P - ((k+1)*0.1)*(P - lo) admitting >= 3 peaks
            // (same rung as "first tau_k <= third-highest peak"); the loop is uniform per window, so its
            // shuffles name only the window's 8 lanes ----
            const bool has = npk >= 3;
            double tau = INFINITY;
            if (has) {
                const double R = __dsub_rn(pm, lo);
                tau = lo;
#pragma unroll 1
                for (int kk = 0; kk < 9; kk++) {
                    const double t = __dsub_rn(pm, __dmul_rn(__dmul_rn((double)(kk + 1), 0.1), R));
                    unsigned mt = 0;
#pragma unroll
                    for (int k = 0; k < KITER; k++) mt |= (v[k] >= t ? 1u : 0u) << k;
                    int c3 = __popc(mt & pk);
#pragma unroll
                    for (int m = 4; m >= 1; m >>= 1) c3 += __shfl_xor_sync(gmask, c3, m);
                    if (c3 >= 3) { tau = t; break; }
                }
            }
            // admitted peaks of this lane as a bit mask over k; i = sub + 8k is monotone in k, so the lane's
            // first / last admitted index and their count come from ffs / clz / popc
            unsigned adm = 0;
#pragma unroll
            for (int k = 0; k < KITER; k++) adm |= (v[k] >= tau ? 1u : 0u) << k;
            adm &= pk;
            int first = adm ? sub + G * (__ffs(adm) - 1) : 0x7fffffff;
            int last = adm ? sub + G * (31 - __clz(adm)) : -1;
            int cnt = __popc(adm);
#pragma unroll
            for (int m = 4; m >= 1; m >>= 1) {
                first = min(first, __shfl_xor_sync(0xffffffffu, first, m));
                last = max(last, __shfl_xor_sync(0xffffffffu, last, m));
                cnt += __shfl_xor_sync(0xffffffffu, cnt, m);
            }
            __syncwarp();                              // all bin increments of this axis are in
            // ---- 14 outputs of this axis: lanes take j = sub and j = sub + 8 ----
            if (live) {
#pragma unroll
                for (int r = 0; r < 2; r++) {
                    const int j = sub + G * r;
                    if (j < 10) {
                        dst[a * 10 + j] = (double)bins_w[j] / dL;
                    } else if (j == 10) {
                        dst[30 + a] = mean;
                    } else if (j == 11) {
                        double pv = qnan;
                        if (has) pv = ((double)(ts[st + last] - ts[st + first]) / (double)(cnt - 1)) / 1.0e6;
                        dst[33 + a] = pv;
                    } else if (j == 12) {
                        dst[36 + a] = ab / dL;
                    } else if (j == 13) {
                        dst[39 + a] = sqrt(sq / dL);
                    }
                }
            }
            __syncwarp();                              // counters are re-zeroed at the top of the next axis
        }
        // ---- RESULTANT: sqrt via fp32 rsqrt seed + one fp64 Newton step (rel. error ~1e-14) ----
        double rs = 0.0;
#pragma unroll
        for (int k = 0; k < KITER; k++) {
            if (EXACT || sub + G * k < L) {
                const double s2 = r2[k * 128];
                const double r = (double)rsqrtf((float)s2);
                double yv = s2 * r;
                yv = fma(0.5 * r, fma(-yv, yv, s2), yv);
                if (!(s2 > 1e-30 && s2 < 1e30)) yv = precise_sqrt(s2);     // out of fp32 range: rare, real branch
                rs += yv;
            }
        }
#pragma unroll
        for (int m = 4; m >= 1; m >>= 1) rs += shfl_xor_d(rs, m);
        if (live && sub == 0) dst[42] = rs / dL;
    }
}
