// Raw WISDM text -> the six stream columns (the ingest ahead of the window transform).
//
// The reference starts from an already transformed CSV (Main/main.py:18-26 reads wisdm_data.csv); the raw release
// it was derived from (WISDM_ar_v1.1_raw.txt) is a sequence of records
//     user,activity,timestamp,x,y,z;
// one or several per line, some of them broken (missing fields, no terminator).  This file parses such text on the
// GPU: record boundaries by a scan over ';', one thread per record, decimal -> double by the exact fast path
// (<= 15 significant digits and |decimal exponent| <= 22: one correctly rounded multiply or divide, Clinger 1990),
// which is what strtod / java.lang.Double.parseDouble return for such literals.  Longer literals are reported,
// not approximated.  Rules (restated in oracle/har_oracle.c: orc_parse_wisdm_raw):
//   * a record is the bytes between ';' terminators with blanks (space, tab, CR, LF) trimmed; empty records vanish;
//   * it must have exactly six non-empty comma-separated fields: int user, activity name (one of the names given by
//     the caller, index = position), int64 timestamp, three decimals; anything else is dropped and counted;
//   * text after the last ';' is one more record (kept if well formed).
#include <cub/cub.cuh>

#include "har_common.cuh"

namespace har {
namespace {

constexpr int ING_CHUNK = 4096;         // bytes per warp in the terminator passes
constexpr int ING_MAX_NAMES = 16, ING_NAME_LEN = 24;

struct ActivityNames { int n; int len[ING_MAX_NAMES]; char s[ING_MAX_NAMES][ING_NAME_LEN]; };

__device__ __forceinline__ bool is_blank(unsigned char c) { return c == ' ' || c == '\t' || c == '\r' || c == '\n'; }

// Terminator passes: a warp owns ING_CHUNK consecutive bytes and walks them in 512-byte rows, one 16-byte load per
// lane (fully coalesced); ';' bytes are found with a SIMD-in-register compare.
__device__ __forceinline__ uint32_t semis16(const unsigned char* __restrict__ text, int64_t nbytes, int64_t off, bool aligned) {
    // bit k of the result: byte off + k is ';'
    uint32_t m = 0;
    if (aligned && off + 16 <= nbytes) {
        const uint4 q = *reinterpret_cast<const uint4*>(text + off);
        const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const uint32_t eq = __vcmpeq4(w[j], 0x3b3b3b3bu);          // 0xff in every byte that is ';'
            m |= ((eq & 1u) | ((eq >> 7) & 2u) | ((eq >> 14) & 4u) | ((eq >> 21) & 8u)) << (4 * j);
        }
    } else {
        for (int k = 0; k < 16 && off + k < nbytes; k++) m |= (text[off + k] == ';' ? 1u : 0u) << k;
    }
    return m;
}

__global__ void k_ing_count(const unsigned char* __restrict__ text, int64_t nbytes, int32_t* __restrict__ cnt) {
    const int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const int64_t b0 = w * ING_CHUNK;
    if (b0 >= nbytes) return;
    const bool aligned = (reinterpret_cast<uintptr_t>(text) & 15) == 0;
    int c = 0;
    for (int it = 0; it < ING_CHUNK / 512; it++) {
        const int64_t off = b0 + it * 512 + lane * 16;
        if (off < nbytes) c += __popc(semis16(text, nbytes, off, aligned));
    }
    c = __reduce_add_sync(0xffffffffu, c);
    if (lane == 0) cnt[w] = c;
}

__global__ void k_ing_ends(const unsigned char* __restrict__ text, int64_t nbytes, const int32_t* __restrict__ off_in,
                           int64_t* __restrict__ rec_end) {
    const int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const int64_t b0 = w * ING_CHUNK;
    if (b0 >= nbytes) return;
    const bool aligned = (reinterpret_cast<uintptr_t>(text) & 15) == 0;
    int64_t k = off_in[w];
    for (int it = 0; it < ING_CHUNK / 512; it++) {
        const int64_t off = b0 + it * 512 + lane * 16;
        uint32_t m = off < nbytes ? semis16(text, nbytes, off, aligned) : 0u;
        const int c = __popc(m);
        int incl = c;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { int o = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += o; }
        int64_t o = k + incl - c;
        while (m) { const int bit = __ffs(m) - 1; rec_end[o++] = off + bit; m &= m - 1; }
        k += __shfl_sync(0xffffffffu, incl, 31);
    }
}

__constant__ double kPow10[23] = {1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11, 1e12, 1e13, 1e14, 1e15, 1e16,
                                  1e17, 1e18, 1e19, 1e20, 1e21, 1e22};

// decimal literal [p, q) -> double; 0 = ok, 1 = malformed, 2 = needs more than the exact fast path
__device__ int parse_decimal(const unsigned char* p, const unsigned char* q, double* out) {
    if (p >= q) return 1;
    bool neg = false;
    if (*p == '-' || *p == '+') { neg = *p == '-'; p++; }
    unsigned long long m = 0;
    int digits = 0, sig = 0, e10 = 0;
    bool any = false, dropped_nonzero = false;
    for (; p < q && *p >= '0' && *p <= '9'; p++) {
        any = true;
        if (sig < 19) { m = m * 10 + (*p - '0'); if (m) sig++; } else { e10++; dropped_nonzero |= *p != '0'; }
        digits++;
    }
    if (p < q && *p == '.') {
        p++;
        for (; p < q && *p >= '0' && *p <= '9'; p++) {
            any = true;
            if (sig < 19) { m = m * 10 + (*p - '0'); if (m) sig++; e10--; } else dropped_nonzero |= *p != '0';
        }
    }
    if (!any) return 1;
    if (p < q && (*p == 'e' || *p == 'E')) {
        p++;
        bool eneg = false;
        if (p < q && (*p == '-' || *p == '+')) { eneg = *p == '-'; p++; }
        if (p >= q) return 1;
        int e = 0;
        for (; p < q && *p >= '0' && *p <= '9'; p++) e = e < 10000 ? e * 10 + (*p - '0') : e;
        e10 += eneg ? -e : e;
    }
    if (p != q) return 1;
    double v;
    if (m == 0) v = 0.0;
    else {
        // strip trailing decimal zeros so that short literals written with padding stay on the fast path
        while (m % 10 == 0 && e10 < 0) { m /= 10; e10++; }
        if (dropped_nonzero || m >= (1ULL << 53) || e10 < -22 || e10 > 22) return 2;
        v = e10 < 0 ? __ddiv_rn((double)m, kPow10[-e10]) : __dmul_rn((double)m, kPow10[e10]);
    }
    *out = neg ? -v : v;
    return 0;
}

__device__ int parse_int64(const unsigned char* p, const unsigned char* q, long long* out) {
    if (p >= q) return 1;
    bool neg = false;
    if (*p == '-' || *p == '+') { neg = *p == '-'; p++; }
    if (p >= q || q - p > 18) return 1;
    long long v = 0;
    for (; p < q; p++) { if (*p < '0' || *p > '9') return 1; v = v * 10 + (*p - '0'); }
    *out = neg ? -v : v;
    return 0;
}

struct Rec { int user, act; long long ts; double x, y, z; };

// 0 = empty (vanishes), 1 = valid, 2 = dropped, 3 = valid shape but a literal outside the exact fast path
__device__ int parse_record(const unsigned char* text, int64_t a, int64_t b, const ActivityNames& names, Rec* r) {
    while (a < b && is_blank(text[a])) a++;
    while (b > a && is_blank(text[b - 1])) b--;
    if (a >= b) return 0;
    // fields are consumed as they are found (no table of field pointers: it would live in local memory)
    const unsigned char* p = text + a; const unsigned char* const e = text + b;
    long long u = 0, ts = 0;
    int act = -1, worst = 0;
    double v[3] = {0.0, 0.0, 0.0};
#pragma unroll
    for (int k = 0; k < 6; k++) {
        const unsigned char* f0 = p;
        while (p < e && *p != ',') p++;
        const unsigned char* f1 = p;
        if (k < 5) { if (p == e) return 2; p++; }       // a comma must follow the first five fields
        else if (p != e) return 2;                      // and none the sixth
        while (f0 < f1 && is_blank(*f0)) f0++;          // blanks around a field are tolerated, empty fields are not
        while (f1 > f0 && is_blank(*(f1 - 1))) f1--;
        if (f0 >= f1) return 2;
        if (k == 0) {
            if (parse_int64(f0, f1, &u) || u < -2147483647LL || u > 2147483647LL) return 2;
        } else if (k == 1) {
            const int alen = (int)(f1 - f0);
            for (int q = 0; q < names.n && act < 0; q++) {
                if (names.len[q] != alen) continue;
                bool same = true;
                for (int i = 0; i < alen && same; i++) same = names.s[q][i] == (char)f0[i];
                if (same) act = q;
            }
            if (act < 0) return 2;
        } else if (k == 2) {
            if (parse_int64(f0, f1, &ts)) return 2;
        } else {
            const int rc = parse_decimal(f0, f1, &v[k - 3]);
            if (rc == 1) return 2;
            worst = max(worst, rc);
        }
    }
    if (worst == 2) return 3;
    r->user = (int)u; r->act = act; r->ts = ts; r->x = v[0]; r->y = v[1]; r->z = v[2];
    return 1;
}

__device__ __forceinline__ void record_span(const int64_t* rec_end, int64_t nrec, int64_t nbytes, int64_t i, int64_t* a, int64_t* b) {
    *a = i ? rec_end[i - 1] + 1 : 0;
    *b = i < nrec ? rec_end[i] : nbytes;               // i == nrec: the unterminated tail
}

// The 128 records of a CTA are contiguous text (about 7 KB): it is copied to shared memory with coalesced loads and
// parsed there.  Parsing straight from global memory made every byte load of a warp touch ~16 different sectors
// (32 records, 58 bytes apart) and the two parse kernels LSU-bound (112 GB/s of text).  Spans longer than the
// staging buffer (garbage without terminators) are parsed from global memory.
constexpr int ING_REC_THREADS = 128, ING_STAGE = 20 * 1024;

__device__ __forceinline__ const unsigned char* stage_records(const unsigned char* __restrict__ text, int64_t nbytes,
                                                              const int64_t* __restrict__ rec_end, int64_t nrec,
                                                              unsigned char* sm, int64_t* base_out) {
    const int64_t i0 = (int64_t)blockIdx.x * ING_REC_THREADS;
    const int64_t i1 = min(i0 + ING_REC_THREADS - 1, nrec);           // last record of the CTA (slot nrec = the tail)
    int64_t a0, b0, a1, b1;
    record_span(rec_end, nrec, nbytes, i0, &a0, &b0);
    record_span(rec_end, nrec, nbytes, i1, &a1, &b1);
    const int64_t len = b1 - a0;
    if (len > ING_STAGE) { *base_out = 0; return text; }              // uniform for the CTA
    for (int64_t k = threadIdx.x; k < len; k += ING_REC_THREADS) sm[k] = text[a0 + k];
    __syncthreads();
    *base_out = a0;
    return sm;
}

__global__ void __launch_bounds__(ING_REC_THREADS) k_ing_classify(const unsigned char* __restrict__ text, int64_t nbytes, const int64_t* __restrict__ rec_end,
                               int64_t nrec, ActivityNames names, int32_t* __restrict__ valid,
                               unsigned long long* __restrict__ counters /* [dropped, unsupported] */) {
    __shared__ __align__(16) unsigned char sm[ING_STAGE];
    int64_t base;
    const unsigned char* src = stage_records(text, nbytes, rec_end, nrec, sm, &base);
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i > nrec) return;
    int64_t a, b;
    record_span(rec_end, nrec, nbytes, i, &a, &b);
    Rec r;
    const int rc = parse_record(src, a - base, b - base, names, &r);
    valid[i] = rc == 1 ? 1 : 0;
    if (rc == 2) atomicAdd(&counters[0], 1ULL);
    if (rc == 3) atomicAdd(&counters[1], 1ULL);
}

__global__ void __launch_bounds__(ING_REC_THREADS) k_ing_emit(const unsigned char* __restrict__ text, int64_t nbytes, const int64_t* __restrict__ rec_end,
                           int64_t nrec, ActivityNames names, const int32_t* __restrict__ valid,
                           const int32_t* __restrict__ pos, int64_t capacity, int32_t* __restrict__ user,
                           int32_t* __restrict__ act, int64_t* __restrict__ ts, double* __restrict__ x,
                           double* __restrict__ y, double* __restrict__ z) {
    __shared__ __align__(16) unsigned char sm[ING_STAGE];
    int64_t base;
    const unsigned char* src = stage_records(text, nbytes, rec_end, nrec, sm, &base);
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i > nrec || !valid[i]) return;
    const int64_t o = pos[i];
    if (o >= capacity) return;
    int64_t a, b;
    record_span(rec_end, nrec, nbytes, i, &a, &b);
    Rec r;
    parse_record(src, a - base, b - base, names, &r);
    user[o] = r.user; act[o] = r.act; ts[o] = r.ts; x[o] = r.x; y[o] = r.y; z[o] = r.z;
}

}  // namespace
}  // namespace har

using namespace har;

HAR_API int har_parse_wisdm_raw(const char* d_text, int64_t nbytes, const char* const* activity_names, int32_t num_names,
                                int64_t capacity, int32_t* d_user, int32_t* d_activity, int64_t* d_ts, double* d_x,
                                double* d_y, double* d_z, int64_t* num_records, int64_t* num_dropped, har_stream_t stream) {
    if (!num_records) return fail(HAR_E_INVALID, "har_parse_wisdm_raw: num_records is NULL");
    *num_records = 0;
    if (num_dropped) *num_dropped = 0;
    if (nbytes < 0 || capacity < 0 || num_names < 1 || num_names > ING_MAX_NAMES || !activity_names)
        return fail(HAR_E_INVALID, "har_parse_wisdm_raw: bad sizes (1..%d activity names)", ING_MAX_NAMES);
    if (nbytes >= (int64_t)ING_CHUNK * 0x7fffffffLL) return fail(HAR_E_UNSUPPORTED, "har_parse_wisdm_raw: text too large for one call");
    ActivityNames names;
    memset(&names, 0, sizeof names);
    names.n = num_names;
    for (int k = 0; k < num_names; k++) {
        const size_t L = activity_names[k] ? strlen(activity_names[k]) : 0;
        if (L == 0 || L >= (size_t)ING_NAME_LEN) return fail(HAR_E_INVALID, "har_parse_wisdm_raw: activity name %d is empty or too long", k);
        names.len[k] = (int)L;
        memcpy(names.s[k], activity_names[k], L);
    }
    if (nbytes == 0) return HAR_OK;
    cudaStream_t s = (cudaStream_t)stream;
    const unsigned char* text = reinterpret_cast<const unsigned char*>(d_text);
    const int64_t nchunks = ceil_div(nbytes, ING_CHUNK);
    DevBuf<int32_t> d_cnt, d_off;
    HAR_CUDA(d_cnt.alloc(nchunks + 1, s)); HAR_CUDA(d_off.alloc(nchunks + 1, s));
    HAR_CUDA(cudaMemsetAsync(d_cnt.p, 0, sizeof(int32_t) * (nchunks + 1), s));
    k_ing_count<<<(unsigned)ceil_div(nchunks * 32, 256), 256, 0, s>>>(text, nbytes, d_cnt.p);
    HAR_LAUNCH_CHECK();
    size_t tmp_bytes = 0;
    HAR_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, d_cnt.p, d_off.p, (int)(nchunks + 1), s));
    DevBuf<unsigned char> d_tmp;
    HAR_CUDA(d_tmp.alloc(tmp_bytes + 16, s));
    HAR_CUDA(cub::DeviceScan::ExclusiveSum(d_tmp.p, tmp_bytes, d_cnt.p, d_off.p, (int)(nchunks + 1), s));
    int32_t nrec32 = 0;
    HAR_CUDA(cudaMemcpyAsync(&nrec32, d_off.p + nchunks, sizeof(int32_t), cudaMemcpyDeviceToHost, s));
    HAR_CUDA(cudaStreamSynchronize(s));
    const int64_t nrec = nrec32;                                     // terminated records; slot nrec is the tail
    DevBuf<int64_t> d_end;
    HAR_CUDA(d_end.alloc(nrec + 1, s));
    k_ing_ends<<<(unsigned)ceil_div(nchunks * 32, 256), 256, 0, s>>>(text, nbytes, d_off.p, d_end.p);
    HAR_LAUNCH_CHECK();
    DevBuf<int32_t> d_valid, d_pos;
    DevBuf<unsigned long long> d_counters;
    HAR_CUDA(d_valid.alloc(nrec + 2, s)); HAR_CUDA(d_pos.alloc(nrec + 2, s)); HAR_CUDA(d_counters.alloc(2, s));
    HAR_CUDA(cudaMemsetAsync(d_counters.p, 0, 2 * sizeof(unsigned long long), s));
    HAR_CUDA(cudaMemsetAsync(d_valid.p, 0, sizeof(int32_t) * (nrec + 2), s));
    const unsigned gb = (unsigned)ceil_div(nrec + 1, ING_REC_THREADS);
    k_ing_classify<<<gb, ING_REC_THREADS, 0, s>>>(text, nbytes, d_end.p, nrec, names, d_valid.p, d_counters.p);
    HAR_LAUNCH_CHECK();
    size_t tmp2 = 0;
    HAR_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp2, d_valid.p, d_pos.p, (int)(nrec + 2), s));
    DevBuf<unsigned char> d_tmp2;
    HAR_CUDA(d_tmp2.alloc(tmp2 + 16, s));
    HAR_CUDA(cub::DeviceScan::ExclusiveSum(d_tmp2.p, tmp2, d_valid.p, d_pos.p, (int)(nrec + 2), s));
    int32_t nvalid = 0;
    unsigned long long counters[2] = {0, 0};
    HAR_CUDA(cudaMemcpyAsync(&nvalid, d_pos.p + nrec + 1, sizeof(int32_t), cudaMemcpyDeviceToHost, s));
    HAR_CUDA(cudaMemcpyAsync(counters, d_counters.p, sizeof counters, cudaMemcpyDeviceToHost, s));
    HAR_CUDA(cudaStreamSynchronize(s));
    if (counters[1])
        return fail(HAR_E_UNSUPPORTED, "har_parse_wisdm_raw: %llu record(s) hold a decimal with more than 15 significant digits or a "
                                       "decimal exponent outside [-22, 22]; only literals on the exact fast path are parsed",
                    counters[1]);
    *num_records = nvalid;
    if (num_dropped) *num_dropped = (int64_t)counters[0];
    if (nvalid > capacity)
        return fail(HAR_E_INVALID, "har_parse_wisdm_raw: %d records do not fit the output capacity %lld", nvalid, (long long)capacity);
    if (nvalid > 0) {
        k_ing_emit<<<gb, ING_REC_THREADS, 0, s>>>(text, nbytes, d_end.p, nrec, names, d_valid.p, d_pos.p, capacity, d_user, d_activity, d_ts,
                                      d_x, d_y, d_z);
        HAR_LAUNCH_CHECK();
    }
    return HAR_OK;
}

// Upper bound on the records of a text (terminators + 1): what to size the output columns with before parsing once.
HAR_API int har_count_wisdm_raw(const char* d_text, int64_t nbytes, int64_t* max_records, har_stream_t stream) {
    if (!max_records) return fail(HAR_E_INVALID, "har_count_wisdm_raw: max_records is NULL");
    *max_records = 0;
    if (nbytes <= 0) return HAR_OK;
    cudaStream_t s = (cudaStream_t)stream;
    const int64_t nchunks = ceil_div(nbytes, ING_CHUNK);
    DevBuf<int32_t> d_cnt;
    DevBuf<long long> d_sum;
    HAR_CUDA(d_cnt.alloc(nchunks, s)); HAR_CUDA(d_sum.alloc(1, s));
    k_ing_count<<<(unsigned)ceil_div(nchunks * 32, 256), 256, 0, s>>>(reinterpret_cast<const unsigned char*>(d_text), nbytes, d_cnt.p);
    HAR_LAUNCH_CHECK();
    size_t tmp_bytes = 0;
    HAR_CUDA(cub::DeviceReduce::Sum(nullptr, tmp_bytes, d_cnt.p, d_sum.p, (int)nchunks, s));
    DevBuf<unsigned char> d_tmp;
    HAR_CUDA(d_tmp.alloc(tmp_bytes + 16, s));
    HAR_CUDA(cub::DeviceReduce::Sum(d_tmp.p, tmp_bytes, d_cnt.p, d_sum.p, (int)nchunks, s));
    long long total = 0;
    HAR_CUDA(cudaMemcpyAsync(&total, d_sum.p, sizeof total, cudaMemcpyDeviceToHost, s));
    HAR_CUDA(cudaStreamSynchronize(s));
    *max_records = (int64_t)total + 1;
    return HAR_OK;
}
