// Synthetic tri-axial accelerometer stream: bench / test input generated in HBM.
// Counter-based (splitmix64 of the global sample index), integer noise scaled by one IEEE multiply,
// carrier from a host-computed 1024-entry sine table: no sin/exp/rand on the device, so the CPU
// oracle (orc_synth_stream) produces bit-identical samples.  Not on the reference's path.
#include "har_common.cuh"

namespace {

__device__ __forceinline__ uint64_t splitmix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

struct SynthTables {
    double mu[18], amp[18], sigma[18];
    int32_t freq[6];
};

__global__ void __launch_bounds__(256)
k_synth(int64_t n, int64_t first_sample, uint64_t seed, const int64_t* __restrict__ bout_start,
        const int32_t* __restrict__ bout_user, const int32_t* __restrict__ bout_act,
        const int64_t* __restrict__ bout_ts0, const double* __restrict__ bout_gain,
        const double* __restrict__ bout_tilt, int64_t num_bouts, const double* __restrict__ sine,
        SynthTables tb, int32_t* __restrict__ user, int32_t* __restrict__ act, int64_t* __restrict__ ts,
        double* __restrict__ x, double* __restrict__ y, double* __restrict__ z) {
    __shared__ double s_sine[1024];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) s_sine[i] = sine[i];
    __syncthreads();
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t g = first_sample + i;
        int64_t lo = 0, hi = num_bouts - 1;          // last bout with bout_start <= g
        while (lo < hi) {
            int64_t mid = (lo + hi + 1) >> 1;
            if (bout_start[mid] <= g) lo = mid; else hi = mid - 1;
        }
        const int64_t k = g - bout_start[lo];
        const int a = bout_act[lo];
        user[i] = bout_user[lo];
        act[i] = a;
        ts[i] = bout_ts0[lo] + k * 50000000LL;
        const double gain = bout_gain[lo], tilt = bout_tilt[lo];
        const double tiltv[3] = {3.0, -4.0, 3.5};
        double v[3];
#pragma unroll
        for (int ax = 0; ax < 3; ax++) {
            uint64_t r = splitmix64(seed ^ ((uint64_t)g * 0x9E3779B97F4A7C15ULL) ^ (uint64_t)(ax + 1));
            long long nz = (long long)(r & 0xffff) + (long long)((r >> 16) & 0xffff) +
                           (long long)((r >> 32) & 0xffff) + (long long)((r >> 48) & 0xffff) - 2 * 65535;
            double noise = __dmul_rn((double)nz, __dmul_rn(tb.sigma[a * 3 + ax], 1.0 / 37837.0));
            double car = __dmul_rn(__dmul_rn(tb.amp[a * 3 + ax], gain), s_sine[(uint64_t)(k * tb.freq[a] + ax * 85) & 1023]);
            v[ax] = __dadd_rn(__dadd_rn(__dadd_rn(tb.mu[a * 3 + ax], __dmul_rn(tilt, tiltv[ax])), car), noise);
        }
        x[i] = v[0]; y[i] = v[1]; z[i] = v[2];
    }
}

}  // namespace

HAR_API int har_synth_stream(int64_t n, int64_t first_sample, uint64_t seed, const int64_t* bout_start,
                             const int32_t* bout_user, const int32_t* bout_activity, const int64_t* bout_ts0,
                             const double* bout_gain, const double* bout_tilt, int64_t num_bouts, const double* sine1024, const double* mu, const double* amp,
                             const double* sigma, const int32_t* freq, int32_t* d_user, int32_t* d_activity,
                             int64_t* d_ts, double* d_x, double* d_y, double* d_z, har_stream_t stream) {
    if (n <= 0) return HAR_OK;
    if (num_bouts <= 0) return har::fail(HAR_E_INVALID, "har_synth_stream: empty bout table");
    cudaStream_t s = (cudaStream_t)stream;
    har::DevBuf<int64_t> bs, bt;
    har::DevBuf<int32_t> bu, ba;
    har::DevBuf<double> sn, bg, bl;
    HAR_CUDA(bs.alloc(num_bouts + 1, s)); HAR_CUDA(bt.alloc(num_bouts, s));
    HAR_CUDA(bu.alloc(num_bouts, s)); HAR_CUDA(ba.alloc(num_bouts, s)); HAR_CUDA(sn.alloc(1024, s));
    HAR_CUDA(cudaMemcpyAsync(bs.p, bout_start, sizeof(int64_t) * (num_bouts + 1), cudaMemcpyHostToDevice, s));
    HAR_CUDA(cudaMemcpyAsync(bt.p, bout_ts0, sizeof(int64_t) * num_bouts, cudaMemcpyHostToDevice, s));
    HAR_CUDA(bg.alloc(num_bouts, s)); HAR_CUDA(bl.alloc(num_bouts, s));
    HAR_CUDA(cudaMemcpyAsync(bg.p, bout_gain, sizeof(double) * num_bouts, cudaMemcpyHostToDevice, s));
    HAR_CUDA(cudaMemcpyAsync(bl.p, bout_tilt, sizeof(double) * num_bouts, cudaMemcpyHostToDevice, s));
    HAR_CUDA(cudaMemcpyAsync(bu.p, bout_user, sizeof(int32_t) * num_bouts, cudaMemcpyHostToDevice, s));
    HAR_CUDA(cudaMemcpyAsync(ba.p, bout_activity, sizeof(int32_t) * num_bouts, cudaMemcpyHostToDevice, s));
    HAR_CUDA(cudaMemcpyAsync(sn.p, sine1024, sizeof(double) * 1024, cudaMemcpyHostToDevice, s));
    SynthTables tb;
    memcpy(tb.mu, mu, sizeof tb.mu); memcpy(tb.amp, amp, sizeof tb.amp); memcpy(tb.sigma, sigma, sizeof tb.sigma);
    memcpy(tb.freq, freq, sizeof tb.freq);
    int64_t blocks = har::ceil_div(n, 256);
    if (blocks > har::kNumSMs * 32) blocks = har::kNumSMs * 32;
    k_synth<<<(unsigned)blocks, 256, 0, s>>>(n, first_sample, seed, bs.p, bu.p, ba.p, bt.p, bg.p, bl.p, num_bouts, sn.p, tb,
                                            d_user, d_activity, d_ts, d_x, d_y, d_z);
    HAR_LAUNCH_CHECK();
    HAR_CUDA(cudaStreamSynchronize(s));     // host tables must outlive the async copies
    return HAR_OK;
}
