// Device-side Matrix Market parse (the text half of the ingest): the token stream the reference reads with
// `ifstream >>` (include/mtx_reader.hpp:17-62: "rows cols nnz", then nnz 1-based "row col value" triples, any whitespace
// between tokens) is cut into 256-byte chunks; one thread per chunk counts the tokens that START in it, an exclusive scan
// gives every chunk the index of its first token, and a second pass parses every token where it lies and stores it by index
// (token 3 + 3 e + f is field f of entry e).  No thread depends on another one's parse.
//
// Decimal -> double is the Eisel-Lemire conversion (D. Lemire, "Number parsing at a gigabyte per second", 2021, restated
// from the paper: a 64 x 128-bit product with a truncated power of five, tools/gen_pow5.py writes the table): exact
// (correctly rounded, ties to even -- what strtod / from_chars of the host reader return) for up to 19 significant digits.
// Tokens it does not decide (more digits, "inf" / "nan", the paper's rare "cannot tell" case) are not guessed: their text
// goes to a list and the HOST converts exactly those with strtod.
//
// The bodies are plain functions over (text, chunk index) so that tests/cpp/mtx_emu.cpp runs the same source on the host
// against Python's float() (millions of tokens, halfway cases, subnormals) before any GPU sees it.
#pragma once
#include <cstdint>

#if defined(__CUDACC__)
#define MP_HD __host__ __device__
#else
#define MP_HD
#endif

namespace spchol {

constexpr int MP_CHUNK = 256;          // bytes per thread
constexpr int MP_TOK_MAX = 64;         // longest token handed to the host (bytes incl. the terminating 0); longer ones are errors

static const unsigned long long mp_pow5_host[651][2] = {
#include "mtx_pow5.inc"
};
#if defined(__CUDACC__)
static __device__ const unsigned long long mp_pow5_dev[651][2] = {
#include "mtx_pow5.inc"
};
#endif

MP_HD inline void mp_pow5(int q, uint64_t& hi, uint64_t& lo) {          // q in [-342, 308]
#if defined(__CUDA_ARCH__)
    hi = mp_pow5_dev[q + 342][0]; lo = mp_pow5_dev[q + 342][1];
#else
    hi = mp_pow5_host[q + 342][0]; lo = mp_pow5_host[q + 342][1];
#endif
}
MP_HD inline void mp_mul64(uint64_t a, uint64_t b, uint64_t& hi, uint64_t& lo) {
#if defined(__CUDA_ARCH__)
    lo = a * b; hi = __umul64hi(a, b);
#else
    const unsigned __int128 p = static_cast<unsigned __int128>(a) * b;
    lo = static_cast<uint64_t>(p); hi = static_cast<uint64_t>(p >> 64);
#endif
}
MP_HD inline int mp_clz64(uint64_t x) {                                 // x != 0
#if defined(__CUDA_ARCH__)
    return __clzll(static_cast<long long>(x));
#else
    return __builtin_clzll(x);
#endif
}
MP_HD inline double mp_from_bits(uint64_t b) {
#if defined(__CUDA_ARCH__)
    return __longlong_as_double(static_cast<long long>(b));
#else
    double d; __builtin_memcpy(&d, &b, 8); return d;
#endif
}

// w * 10^q (w != 0) as the bits of the nearest double (ties to even).  Returns 0, or 1 when the truncated product cannot
// decide the rounding (then nothing is written).
MP_HD inline int mp_eisel_lemire(uint64_t w, int q, uint64_t* bits) {
    if (q < -342) { *bits = 0; return 0; }                               // below half the smallest subnormal for every 64-bit w
    if (q > 308) { *bits = 0x7ffull << 52; return 0; }
    const int lz = mp_clz64(w);
    w <<= lz;
    uint64_t p_hi, p_lo, hi, lo;
    mp_pow5(q, p_hi, p_lo);
    mp_mul64(w, p_hi, hi, lo);
    if ((hi & 0x1ffull) == 0x1ffull) {                                   // the 55 leading bits may still change: second word
        uint64_t s_hi, s_lo;
        mp_mul64(w, p_lo, s_hi, s_lo);
        lo += s_hi;
        if (s_hi > lo) ++hi;
    }
    if (lo == ~0ull && (q < -27 || q > 55)) return 1;                    // the paper's "cannot tell": the host converts it
    const int upperbit = static_cast<int>(hi >> 63);
    const int shift = upperbit + 64 - 52 - 3;
    uint64_t m = hi >> shift;
    int p2 = (((152170 + 65536) * q) >> 16) + 63 + upperbit - lz + 1023;
    if (p2 <= 0) {                                                       // subnormal (or zero)
        if (-p2 + 1 >= 64) { *bits = 0; return 0; }
        m >>= -p2 + 1;
        m += m & 1;
        m >>= 1;
        p2 = m < (1ull << 52) ? 0 : 1;
        *bits = (static_cast<uint64_t>(p2) << 52) | (m & ~(1ull << 52));
        return 0;
    }
    // exactly half way between two doubles: only possible when 5^q fits one word; then round to even instead of up
    if (lo <= 1 && q >= -4 && q <= 23 && (m & 3) == 1 && (m << shift) == hi) m &= ~1ull;
    m += m & 1;
    m >>= 1;
    if (m >= (2ull << 52)) { m = 1ull << 52; ++p2; }
    m &= ~(1ull << 52);
    if (p2 >= 0x7ff) { *bits = 0x7ffull << 52; return 0; }
    *bits = (static_cast<uint64_t>(p2) << 52) | m;
    return 0;
}

MP_HD inline bool mp_ws(char c) { return c == ' ' || c == '\n' || c == '\t' || c == '\r' || c == '\f' || c == '\v'; }
MP_HD inline bool mp_digit(char c) { return c >= '0' && c <= '9'; }

// [+|-] digits [. digits] [(e|E) [+|-] digits] filling the whole token [p, e).  0: *out written; 1: the host must convert this
// token (more than 19 significant digits, inf / nan spellings, undecided rounding); 2: not a number
MP_HD inline int mp_parse_real(const char* p, const char* e, double* out) {
    if (p >= e) return 2;
    bool neg = false;
    if (*p == '+' || *p == '-') { neg = *p == '-'; ++p; }
    if (p < e && ((*p >= 'a' && *p <= 'z') || (*p >= 'A' && *p <= 'Z'))) return 1;         // inf, nan, infinity, ...: strtod knows them
    uint64_t w = 0;
    int nd = 0, exp10 = 0;
    bool any = false, too_long = false;
    for (; p < e && mp_digit(*p); ++p) {
        any = true;
        const int d = *p - '0';
        if (w == 0 && d == 0) continue;                                  // leading zeros
        if (nd < 19) { w = w * 10 + d; ++nd; } else too_long = true;
    }
    if (p < e && *p == '.') {
        ++p;
        for (; p < e && mp_digit(*p); ++p) {
            any = true;
            const int d = *p - '0';
            if (w == 0 && d == 0) { --exp10; continue; }
            if (nd < 19) { w = w * 10 + d; ++nd; --exp10; } else too_long = true;
        }
    }
    if (!any) return 2;
    if (p < e && (*p == 'e' || *p == 'E')) {
        ++p;
        bool eneg = false;
        if (p < e && (*p == '+' || *p == '-')) { eneg = *p == '-'; ++p; }
        if (p >= e || !mp_digit(*p)) return 2;
        int ex = 0;
        for (; p < e && mp_digit(*p); ++p) if (ex < 100000) ex = ex * 10 + (*p - '0');
        exp10 += eneg ? -ex : ex;
    }
    if (p != e) return 2;
    if (too_long) return 1;
    uint64_t bits = 0;
    if (w != 0 && mp_eisel_lemire(w, exp10, &bits)) return 1;
    if (neg) bits |= 1ull << 63;
    *out = mp_from_bits(bits);
    return 0;
}
// [+|-] digits filling the whole token, at most 18 digits.  false: not an integer
MP_HD inline bool mp_parse_int(const char* p, const char* e, long long* out) {
    if (p >= e) return false;
    bool neg = false;
    if (*p == '+' || *p == '-') { neg = *p == '-'; ++p; }
    if (p >= e || e - p > 18) return false;
    long long v = 0;
    for (; p < e; ++p) { if (!mp_digit(*p)) return false; v = v * 10 + (*p - '0'); }
    *out = neg ? -v : v;
    return true;
}

#if defined(__CUDA_ARCH__)
#define MP_ATOMIC_MIN_U64(ptr, v) atomicMin(ptr, v)
#define MP_ATOMIC_INC_U32(ptr) atomicAdd(ptr, 1u)
#else
#define MP_ATOMIC_MIN_U64(ptr, v) do { if ((v) < *(ptr)) *(ptr) = (v); } while (0)
#define MP_ATOMIC_INC_U32(ptr) ((*(ptr))++)
#endif

struct MpOut {
    long long* hdr;                    // [3]: rows, cols, nnz as written in the file
    int* ti; int* tj; double* tx;      // [nz_cap]: 0-based row / column and value of every entry, in file order (not yet mirrored)
    int64_t nz_cap;
    unsigned long long* err_pos;       // smallest index of a token that is not what its place asks for (start: ~0)
    unsigned long long* ntok;          // tokens in the text (written by the last chunk)
    unsigned* fb_count; unsigned fb_cap;
    int* fb_e; char* fb_text;          // values the host converts: entry index and the token (0-terminated, MP_TOK_MAX bytes each)
};

// pass 1: tokens that start in this chunk
MP_HD inline unsigned mp_count_chunk(const char* text, int64_t nbytes, int64_t chunk) {
    const int64_t b0 = chunk * MP_CHUNK, b1 = b0 + MP_CHUNK < nbytes ? b0 + MP_CHUNK : nbytes;
    bool prev_ws = b0 == 0 ? true : mp_ws(text[b0 - 1]);
    unsigned c = 0;
    for (int64_t i = b0; i < b1; ++i) {
        const bool w = mp_ws(text[i]);
        if (!w && prev_ws) ++c;
        prev_ws = w;
    }
    return c;
}
// pass 2: parse them (tok0 = index of the chunk's first token, from the scan of the counts)
MP_HD inline void mp_parse_chunk(const char* text, int64_t nbytes, int64_t chunk, int64_t nchunks, unsigned tok0, const MpOut& o) {
    const int64_t b0 = chunk * MP_CHUNK, b1 = b0 + MP_CHUNK < nbytes ? b0 + MP_CHUNK : nbytes;
    bool prev_ws = b0 == 0 ? true : mp_ws(text[b0 - 1]);
    unsigned long long t = tok0;
    for (int64_t i = b0; i < b1; ++i) {
        const bool w = mp_ws(text[i]);
        const bool start = !w && prev_ws;
        prev_ws = w;
        if (!start) continue;
        int64_t j = i + 1;
        while (j < nbytes && !mp_ws(text[j])) ++j;                       // the token may run into the next chunks
        const char* p = text + i; const char* e = text + j;
        const unsigned long long idx = t++;
        if (idx < 3) {
            long long v;
            if (mp_parse_int(p, e, &v)) o.hdr[idx] = v; else MP_ATOMIC_MIN_U64(o.err_pos, idx);
            continue;
        }
        const unsigned long long en = (idx - 3) / 3;
        const int f = static_cast<int>((idx - 3) % 3);
        if (en >= static_cast<unsigned long long>(o.nz_cap)) continue;    // beyond the entries the caller asked for: never read (as the reference)
        if (f < 2) {
            long long v;
            if (mp_parse_int(p, e, &v) && v - 1 >= -2147483647LL - 1 && v - 1 <= 2147483647LL) (f == 0 ? o.ti : o.tj)[en] = static_cast<int>(v - 1);
            else MP_ATOMIC_MIN_U64(o.err_pos, idx);
        } else {
            double v = 0.0;
            const int st = mp_parse_real(p, e, &v);
            if (st == 0) o.tx[en] = v;
            else if (st == 1 && j - i < MP_TOK_MAX) {
                const unsigned k = MP_ATOMIC_INC_U32(o.fb_count);
                if (k < o.fb_cap) {
                    o.fb_e[k] = static_cast<int>(en);
                    char* dst = o.fb_text + static_cast<int64_t>(k) * MP_TOK_MAX;
                    for (int64_t q = 0; q < j - i; ++q) dst[q] = p[q];
                    dst[j - i] = 0;
                }
            } else MP_ATOMIC_MIN_U64(o.err_pos, idx);
        }
    }
    if (chunk == nchunks - 1) *o.ntok = t;
}
// the reference mirrors every entry into the upper triangle as it reads it (mtx_reader.hpp:50-51)
MP_HD inline void mp_mirror_entry(int64_t e, int* ti, int* tj) {
    const int r = ti[e], c = tj[e];
    if (c < r) { ti[e] = c; tj[e] = r; }
}

#if defined(__CUDACC__)
__global__ void __launch_bounds__(128) k_mp_count(const char* __restrict__ text, int64_t nbytes, int64_t nchunks, unsigned* __restrict__ counts) {
    const int64_t c = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
    if (c < nchunks) counts[c] = mp_count_chunk(text, nbytes, c);
}
__global__ void __launch_bounds__(128) k_mp_parse(const char* __restrict__ text, int64_t nbytes, int64_t nchunks, const unsigned* __restrict__ tok0, MpOut o) {
    const int64_t c = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
    if (c < nchunks) mp_parse_chunk(text, nbytes, c, nchunks, tok0[c], o);
}
__global__ void k_mp_mirror(int64_t nz, int* __restrict__ ti, int* __restrict__ tj) {
    const int64_t e = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
    if (e < nz) mp_mirror_entry(e, ti, tj);
}
__global__ void k_mp_scatter(unsigned n, const int* __restrict__ e, const double* __restrict__ v, double* __restrict__ tx) {
    const unsigned k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n) tx[e[k]] = v[k];
}
#endif

}  // namespace spchol
