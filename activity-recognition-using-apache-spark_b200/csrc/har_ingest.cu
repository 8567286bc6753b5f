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

constexpr int ING_CHUNK = 128;          // bytes per thread in the terminator passes
constexpr int ING_MAX_NAMES = 16, ING_NAME_LEN = 24;

struct ActivityNames { int n; int len[ING_MAX_NAMES]; char s[ING_MAX_NAMES][ING_NAME_LEN]; };

__device__ __forceinline__ bool is_blank(unsigned char c) { return c == ' ' || c == '\t' || c == '\r' || c == '\n'; }

__global__ void k_ing_count(const unsigned char* __restrict__ text, int64_t nbytes, int32_t* __restrict__ cnt) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t b0 = t * ING_CHUNK;
    if (b0 >= nbytes) return;
    const int64_t b1 = min(nbytes, b0 + ING_CHUNK);
    int c = 0;
    for (int64_t i = b0; i < b1; i++) c += text[i] == ';';
    cnt[t] = c;
}

__global__ void k_ing_ends(const unsigned char* __restrict__ text, int64_t nbytes, const int32_t* __restrict__ off,
                           int64_t* __restrict__ rec_end) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t b0 = t * ING_CHUNK;
    if (b0 >= nbytes) return;
    const int64_t b1 = min(nbytes, b0 + ING_CHUNK);
    int64_t k = off[t];
    for (int64_t i = b0; i < b1; i++) if (text[i] == ';') rec_end[k++] = i;
}

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
        const double P10[23] = {1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11, 1e12, 1e13, 1e14, 1e15, 1e16,
                                1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
        v = e10 < 0 ? __ddiv_rn((double)m, P10[-e10]) : __dmul_rn((double)m, P10[e10]);
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
    const unsigned char* f0[6]; const unsigned char* f1[6];
    int nf = 0;
    const unsigned char* p = text + a; const unsigned char* e = text + b;
    const unsigned char* s = p;
    for (;; p++) {
        if (p == e || *p == ',') {
            if (nf == 6) return 2;
            f0[nf] = s; f1[nf] = p; nf++;
            if (p == e) break;
            s = p + 1;
        }
    }
    if (nf != 6) return 2;
    for (int k = 0; k < 6; k++) {                       // blanks around a field are tolerated, empty fields are not
        while (f0[k] < f1[k] && is_blank(*f0[k])) f0[k]++;
        while (f1[k] > f0[k] && is_blank(*(f1[k] - 1))) f1[k]--;
        if (f0[k] >= f1[k]) return 2;
    }
    long long u;
    if (parse_int64(f0[0], f1[0], &u) || u < -2147483647LL || u > 2147483647LL) return 2;
    int act = -1;
    const int alen = (int)(f1[1] - f0[1]);
    for (int k = 0; k < names.n && act < 0; k++) {
        if (names.len[k] != alen) continue;
        bool same = true;
        for (int i = 0; i < alen && same; i++) same = names.s[k][i] == (char)f0[1][i];
        if (same) act = k;
    }
    if (act < 0) return 2;
    long long ts;
    if (parse_int64(f0[2], f1[2], &ts)) return 2;
    double v[3];
    int worst = 0;
    for (int k = 0; k < 3; k++) {
        const int rc = parse_decimal(f0[3 + k], f1[3 + k], &v[k]);
        if (rc == 1) return 2;
        worst = max(worst, rc);
    }
    if (worst == 2) return 3;
    r->user = (int)u; r->act = act; r->ts = ts; r->x = v[0]; r->y = v[1]; r->z = v[2];
    return 1;
}

__device__ __forceinline__ void record_span(const int64_t* rec_end, int64_t nrec, int64_t nbytes, int64_t i, int64_t* a, int64_t* b) {
    *a = i ? rec_end[i - 1] + 1 : 0;
    *b = i < nrec ? rec_end[i] : nbytes;               // i == nrec: the unterminated tail
}

__global__ void k_ing_classify(const unsigned char* __restrict__ text, int64_t nbytes, const int64_t* __restrict__ rec_end,
                               int64_t nrec, ActivityNames names, int32_t* __restrict__ valid,
                               unsigned long long* __restrict__ counters /* [dropped, unsupported] */) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i > nrec) return;
    int64_t a, b;
    record_span(rec_end, nrec, nbytes, i, &a, &b);
    Rec r;
    const int rc = parse_record(text, a, b, names, &r);
    valid[i] = rc == 1 ? 1 : 0;
    if (rc == 2) atomicAdd(&counters[0], 1ULL);
    if (rc == 3) atomicAdd(&counters[1], 1ULL);
}

__global__ void k_ing_emit(const unsigned char* __restrict__ text, int64_t nbytes, const int64_t* __restrict__ rec_end,
                           int64_t nrec, ActivityNames names, const int32_t* __restrict__ valid,
                           const int32_t* __restrict__ pos, int64_t capacity, int32_t* __restrict__ user,
                           int32_t* __restrict__ act, int64_t* __restrict__ ts, double* __restrict__ x,
                           double* __restrict__ y, double* __restrict__ z) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i > nrec || !valid[i]) return;
    const int64_t o = pos[i];
    if (o >= capacity) return;
    int64_t a, b;
    record_span(rec_end, nrec, nbytes, i, &a, &b);
    Rec r;
    parse_record(text, a, b, names, &r);
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
    k_ing_count<<<(unsigned)ceil_div(nchunks, 256), 256, 0, s>>>(text, nbytes, d_cnt.p);
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
    k_ing_ends<<<(unsigned)ceil_div(nchunks, 256), 256, 0, s>>>(text, nbytes, d_off.p, d_end.p);
    HAR_LAUNCH_CHECK();
    DevBuf<int32_t> d_valid, d_pos;
    DevBuf<unsigned long long> d_counters;
    HAR_CUDA(d_valid.alloc(nrec + 2, s)); HAR_CUDA(d_pos.alloc(nrec + 2, s)); HAR_CUDA(d_counters.alloc(2, s));
    HAR_CUDA(cudaMemsetAsync(d_counters.p, 0, 2 * sizeof(unsigned long long), s));
    HAR_CUDA(cudaMemsetAsync(d_valid.p, 0, sizeof(int32_t) * (nrec + 2), s));
    const unsigned gb = (unsigned)ceil_div(nrec + 1, 128);
    k_ing_classify<<<gb, 128, 0, s>>>(text, nbytes, d_end.p, nrec, names, d_valid.p, d_counters.p);
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
        k_ing_emit<<<gb, 128, 0, s>>>(text, nbytes, d_end.p, nrec, names, d_valid.p, d_pos.p, capacity, d_user, d_activity, d_ts,
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
    k_ing_count<<<(unsigned)ceil_div(nchunks, 256), 256, 0, s>>>(reinterpret_cast<const unsigned char*>(d_text), nbytes, d_cnt.p);
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
