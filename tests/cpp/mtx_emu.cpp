// Runs the SOURCE of the device-side Matrix Market parse (csrc/mtx_parse.cuh) on the host, chunk after chunk, so that the
// tokenisation and the decimal -> double conversion are checked (tests/test_mtx_parse_emu.py, against Python's float() and the
// host reader of the product) without a GPU.
#include <cstdint>
#include <cstring>
#include <vector>

#include "mtx_parse.cuh"

using namespace spchol;

// one token per line in `text` (n tokens, offsets off[0..n]): status[k] = mp_parse_real's return, val[k] = its value
extern "C" void mtxemu_reals(const char* text, const int64_t* off, int64_t n, int* status, double* val) {
    for (int64_t k = 0; k < n; ++k) { val[k] = 0.0; status[k] = mp_parse_real(text + off[k], text + off[k + 1] - 1, &val[k]); }
}
// the two passes over a text (what spchol_mtx_parse_device launches), then the mirror; returns the number of chunks
extern "C" int64_t mtxemu_parse(const char* text, int64_t nbytes, int64_t nz_cap, int* ti, int* tj, double* tx, long long* hdr,
                                unsigned long long* err_pos, unsigned long long* ntok, unsigned* fb_count, unsigned fb_cap, int* fb_e,
                                char* fb_text, int mirror) {
    const int64_t nchunks = (nbytes + MP_CHUNK - 1) / MP_CHUNK;
    std::vector<unsigned> cnt(static_cast<size_t>(nchunks) + 1, 0);
    for (int64_t c = 0; c < nchunks; ++c) cnt[c] = mp_count_chunk(text, nbytes, c);
    unsigned run = 0;
    for (int64_t c = 0; c < nchunks; ++c) { const unsigned v = cnt[c]; cnt[c] = run; run += v; }      // exclusive scan
    *err_pos = ~0ull; *ntok = 0; *fb_count = 0;
    MpOut o{hdr, ti, tj, tx, nz_cap, err_pos, ntok, fb_count, fb_cap, fb_e, fb_text};
    // chunks in reverse order: no chunk may depend on an earlier one having run
    for (int64_t c = nchunks - 1; c >= 0; --c) mp_parse_chunk(text, nbytes, c, nchunks, cnt[c], o);
    if (mirror) for (int64_t e = 0; e < nz_cap; ++e) mp_mirror_entry(e, ti, tj);
    return nchunks;
}
