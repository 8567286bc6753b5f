// k_window_features_v2 — register-resident variant of the window-feature kernel (K2/K3).
//
// ncu on the TMA-staged kernel (profiles/r01_window_features_*.txt) showed it issue-bound, not
// memory-bound: ~6500 warp instructions per window, most of them the 32-lane butterfly reductions and a
// lane-0 epilogue full of fp64 divisions, while DRAM traffic already equalled the algorithmic bytes.
// This version keeps the HBM behaviour (every sample read exactly once, 128 B coalesced per window and
// load) and cuts the instruction count:
//   * 16 lanes per window, two windows per warp: reductions are 4 butterfly steps shared by two windows;
//   * each lane keeps its <= 16 samples of all three axes in registers, so the second pass (deviations,
//     bins, peaks) re-reads nothing;
//   * the 43 outputs are produced lane-parallel (one division per lane per round) instead of by lane 0;
//   * sqrt for RESULTANT = fp32 rsqrt seed + one fp64 Newton step (relative error ~1e-14).
// Discrete decisions (bin index, peak selection, threshold ladder) use exactly the oracle's operations.
#pragma once

template <int KITER>
__global__ void __launch_bounds__(128, 3)
k_window_features_v2(const double* __restrict__ x, const double* __restrict__ y, const double* __restrict__ z,
                     const int64_t* __restrict__ ts, const int64_t* __restrict__ win_start, int64_t nw, int L,
                     double* __restrict__ out, int64_t ld_out) {
    constexpr int G = 16;
    const int lane = threadIdx.x & 31, sub = lane & (G - 1), half = lane >> 4;
    const int64_t pair0 = ((int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5));
    const int64_t npairs = (int64_t)gridDim.x * (blockDim.x >> 5);
    const double* axes[3] = {x, y, z};
    const double dL = (double)L;

    for (int64_t pr = pair0; pr * 2 < nw; pr += npairs) {
        const int64_t w = pr * 2 + half;
        const bool live = w < nw;
        const int64_t st = live ? win_start[w] : win_start[nw - 1];
        // ---- load: lane `sub` owns samples sub, sub+16, ... of each axis ----
        double v[3][KITER];
#pragma unroll
        for (int a = 0; a < 3; a++) {
            const double* p = axes[a] + st;
#pragma unroll
            for (int k = 0; k < KITER; k++) {
                const int i = sub + G * k;
                v[a][k] = i < L ? __ldg(p + i) : 0.0;
            }
        }
        // ---- pass A: sum / min / max per axis, resultant ----
        double sum[3], lo[3], hi[3], rs = 0.0;
#pragma unroll
        for (int a = 0; a < 3; a++) { sum[a] = 0.0; lo[a] = INFINITY; hi[a] = -INFINITY; }
#pragma unroll
        for (int k = 0; k < KITER; k++) {
            if (sub + G * k < L) {
#pragma unroll
                for (int a = 0; a < 3; a++) { sum[a] += v[a][k]; lo[a] = fmin(lo[a], v[a][k]); hi[a] = fmax(hi[a], v[a][k]); }
                const double s2 = fma(v[2][k], v[2][k], fma(v[1][k], v[1][k], v[0][k] * v[0][k]));
                // sqrt(s2): fp32 rsqrt seed, one fp64 Newton step on y = s2 * r
                const float rf = rsqrtf((float)s2);
                const double r = (double)rf;
                double yv = s2 * r;
                yv = fma(0.5 * r, fma(-yv, yv, s2), yv);
                rs += (s2 > 1e-30 && s2 < 1e30) ? yv : sqrt(s2);
            }
        }
#pragma unroll
        for (int m = 8; m >= 1; m >>= 1) {
#pragma unroll
            for (int a = 0; a < 3; a++) {
                sum[a] += shfl_xor_d(sum[a], m);
                lo[a] = fmin(lo[a], shfl_xor_d(lo[a], m));
                hi[a] = fmax(hi[a], shfl_xor_d(hi[a], m));
            }
            rs += shfl_xor_d(rs, m);
        }
        // mean and bin scale: lanes 0..2 of each window do one axis each (one division sequence per warp)
        double mean[3], scale[3];
        {
            const double s_sel = sub == 0 ? sum[0] : (sub == 1 ? sum[1] : sum[2]);
            const double l_sel = sub == 0 ? lo[0] : (sub == 1 ? lo[1] : lo[2]);
            const double h_sel = sub == 0 ? hi[0] : (sub == 1 ? hi[1] : hi[2]);
            const double mean_sel = s_sel / dL;
            const double scale_sel = h_sel > l_sel ? 10.0 / (h_sel - l_sel) : 0.0;
#pragma unroll
            for (int a = 0; a < 3; a++) {
                mean[a] = __shfl_sync(0xffffffffu, mean_sel, half * G + a);
                scale[a] = __shfl_sync(0xffffffffu, scale_sel, half * G + a);
            }
        }
        // ---- pass B: deviations, bins, local maxima ----
        double sq[3], ab[3], p1[3], p2[3], p3[3];
        unsigned long long lo5[3], hi5[3];
        unsigned pk[3];
        int npk[3];
#pragma unroll
        for (int a = 0; a < 3; a++) {
            sq[a] = 0.0; ab[a] = 0.0; lo5[a] = 0; hi5[a] = 0; pk[a] = 0; npk[a] = 0;
            p1[a] = p2[a] = p3[a] = -INFINITY;
            const double* p = axes[a] + st;
#pragma unroll
            for (int k = 0; k < KITER; k++) {
                const int i = sub + G * k;
                if (i < L) {
                    const double c = v[a][k];
                    const double d = c - mean[a];
                    sq[a] = fma(d, d, sq[a]);
                    ab[a] += fabs(d);
                    const int bq = __double2int_rd((c - lo[a]) * scale[a]);
                    const int b = bq >= 9 ? 9 : bq;
                    if (b < 5) lo5[a] += 1ULL << (12 * b); else hi5[a] += 1ULL << (12 * (b - 5));
                    if (i >= 1 && i + 1 < L) {
                        // neighbours belong to the adjacent lanes' lines, just loaded: L1 hits
                        if (c > __ldg(p + i - 1) && c >= __ldg(p + i + 1)) {
                            pk[a] |= 1u << k; npk[a]++;
                            top3_insert(c, p1[a], p2[a], p3[a]);
                        }
                    }
                }
            }
        }
#pragma unroll
        for (int m = 8; m >= 1; m >>= 1) {
#pragma unroll
            for (int a = 0; a < 3; a++) {
                sq[a] += shfl_xor_d(sq[a], m);
                ab[a] += shfl_xor_d(ab[a], m);
                lo5[a] += __shfl_xor_sync(0xffffffffu, lo5[a], m);
                hi5[a] += __shfl_xor_sync(0xffffffffu, hi5[a], m);
                npk[a] += __shfl_xor_sync(0xffffffffu, npk[a], m);
                const double q1 = shfl_xor_d(p1[a], m), q2 = shfl_xor_d(p2[a], m), q3 = shfl_xor_d(p3[a], m);
                top3_insert(q1, p1[a], p2[a], p3[a]); top3_insert(q2, p1[a], p2[a], p3[a]); top3_insert(q3, p1[a], p2[a], p3[a]);
            }
        }
        // ---- peaks: threshold ladder, then first / last / count of the admitted peaks ----
        double dt[3];
        int pc[3];
#pragma unroll
        for (int a = 0; a < 3; a++) {
            dt[a] = 0.0; pc[a] = 0;
            const bool has = npk[a] >= 3;          // uniform within the window's 16 lanes (not across the warp)
            double tau = INFINITY;
            if (has) {
                const double P = p1[a], R = __dsub_rn(P, lo[a]);
                tau = lo[a];
#pragma unroll 1
                for (int k = 0; k < 9; k++) {
                    const double t = __dsub_rn(P, __dmul_rn(__dmul_rn((double)(k + 1), 0.1), R));
                    if (t <= p3[a]) { tau = t; break; }
                }
            }
            int first = 0x7fffffff, last = -1, cnt = 0;
#pragma unroll
            for (int k = 0; k < KITER; k++)
                if (((pk[a] >> k) & 1u) && v[a][k] >= tau) { const int i = sub + G * k; first = min(first, i); last = max(last, i); cnt++; }
#pragma unroll
            for (int m = 8; m >= 1; m >>= 1) {
                first = min(first, __shfl_xor_sync(0xffffffffu, first, m));
                last = max(last, __shfl_xor_sync(0xffffffffu, last, m));
                cnt += __shfl_xor_sync(0xffffffffu, cnt, m);
            }
            if (has) {
                pc[a] = cnt;
                dt[a] = (double)(ts[st + last] - ts[st + first]);
            }
        }
        // ---- outputs: lane `sub` writes outputs sub, sub+16, sub+32 (one division per lane per round) ----
        if (live) {
            double* dst = out + w * ld_out;
#pragma unroll
            for (int r = 0; r < 3; r++) {
                const int o = sub + G * r;
                if (o < 30) {
                    const int a = o / 10, b = o - a * 10;
                    const unsigned long long l5 = a == 0 ? lo5[0] : (a == 1 ? lo5[1] : lo5[2]);
                    const unsigned long long h5 = a == 0 ? hi5[0] : (a == 1 ? hi5[1] : hi5[2]);
                    const unsigned cnt = b < 5 ? (unsigned)((l5 >> (12 * b)) & 0xfff) : (unsigned)((h5 >> (12 * (b - 5))) & 0xfff);
                    dst[o] = (double)cnt / dL;
                } else if (o < 33) {
                    const int a = o - 30;
                    dst[o] = a == 0 ? mean[0] : (a == 1 ? mean[1] : mean[2]);
                } else if (o < 36) {
                    const int a = o - 33;
                    const double num = a == 0 ? dt[0] : (a == 1 ? dt[1] : dt[2]);
                    const int c = a == 0 ? pc[0] : (a == 1 ? pc[1] : pc[2]);
                    dst[o] = c >= 3 ? (num / (double)(c - 1)) / 1.0e6 : __longlong_as_double(0x7ff8000000000000LL);
                } else if (o < 39) {
                    const int a = o - 36;
                    dst[o] = (a == 0 ? ab[0] : (a == 1 ? ab[1] : ab[2])) / dL;
                } else if (o < 42) {
                    const int a = o - 39;
                    dst[o] = sqrt((a == 0 ? sq[0] : (a == 1 ? sq[1] : sq[2])) / dL);
                } else if (o == 42) {
                    dst[o] = rs / dL;
                }
            }
        }
    }
}
