// mtx_reader.hpp -- drop-in for the reference's include/mtx_reader.hpp:17-62.
// Parsing happens in libspchol.so (spchol_mtx_read) with the reference's rules.
#pragma once
#include <stdexcept>
#include <string>
#include <vector>

#include "chol.hpp"

template <typename T>
csc_matrix<T, sym::upper> load_matrix_market_to_csc(const std::string& filename) {
    int n = 0; std::int64_t nz = 0;
    if (spchol_mtx_read(filename.c_str(), &n, &nz, nullptr, nullptr, nullptr) != SPCHOL_OK)
        throw std::runtime_error(spchol_last_error());          // "Could not open file <name>" (reference :20)
    std::vector<int> ti(nz), tj(nz);
    std::vector<T> tx(nz);
    if (spchol_mtx_read(filename.c_str(), &n, &nz, ti.data(), tj.data(), tx.data()) != SPCHOL_OK)
        throw std::runtime_error(spchol_last_error());
    return triplet_to_csc_matrix(ti, tj, tx, n);
}
