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

// The same with both halves on the GPU (beyond the reference): the text parse (spchol_mtx_read_gpu: one thread per 256-byte
// chunk of the file, exact decimal -> double) and the COO -> CSC conversion (spchol_triplet_to_csc_gpu: stable radix sort +
// segmented sum).  Same matrix as load_matrix_market_to_csc, bit for bit.
template <typename T>
csc_matrix<T, sym::upper> load_matrix_market_to_csc_gpu(const std::string& filename) {
    static_assert(std::is_same_v<T, double>, "the B200 path is fp64 (the only instantiation the reference uses)");
    int n = 0; std::int64_t nz = 0;
    if (spchol_mtx_read_gpu(filename.c_str(), &n, &nz, nullptr, nullptr, nullptr, nullptr) != SPCHOL_OK)
        throw std::runtime_error(spchol_last_error());
    std::vector<int> ti(nz), tj(nz);
    std::vector<double> tx(nz);
    if (spchol_mtx_read_gpu(filename.c_str(), &n, &nz, ti.data(), tj.data(), tx.data(), nullptr) != SPCHOL_OK)
        throw std::runtime_error(spchol_last_error());
    std::vector<int> Ap(n + 1), Ai(std::max<std::int64_t>(nz, 1));
    std::vector<double> Ax(std::max<std::int64_t>(nz, 1));
    std::int64_t m = 0;
    if (spchol_triplet_to_csc_gpu(nz, ti.data(), tj.data(), tx.data(), n, Ap.data(), Ai.data(), Ax.data(), &m, nullptr) != SPCHOL_OK)
        throw std::runtime_error(spchol_last_error());
    Ai.resize(m); Ax.resize(m);
    csc_matrix<T, sym::upper> A(n, std::max<std::int64_t>(m, 1));
    A.p() = std::move(Ap); A.i() = std::move(Ai); A.x() = std::move(Ax);
    return A;
}
