// TEST INFRASTRUCTURE — optional BLAS/LAPACK backend of the CPU oracle, for the TIMED CPU baseline only (bench.py).
//
// The reference does its dense linear algebra through Armadillo -> whatever BLAS/LAPACK R links (src/Makevars.in:3):
// arma::chol -> dpotrf (src/cholesky-cache.cpp:23, src/draw-fstar.cpp:36), solve(trimatl(L), X) -> dtrtrs
// (src/cholesky-cache.cpp:6-7, src/draw-fstar.cpp:30), L * z and kstar^T alpha -> dgemv (src/mvnormal.h:23,
// src/draw-fstar.cpp:52), tmp^T tmp -> dsyrk/dgemm (src/draw-fstar.cpp:34). The oracle's own loops (orc_math.h) are the
// arithmetic the parity tests pin; a baseline timed on them would flatter the GPU, so the bench can switch these calls
// to an optimised BLAS: the OpenBLAS that ships inside the scipy wheel of this image (LP64, symbols prefixed scipy_),
// loaded at run time with dlopen — nothing links against it, and the parity tests never turn it on.
// Threading as a BLAS-linked R would see it: calls made from the serial part of a draw_* (dpotrf, dtrsm, dsyrk) use all
// threads of the BLAS; calls made inside the `omp parallel for` over items (dgemv, dtrsv) run single-threaded.
#pragma once
#include <dlfcn.h>

#include <cstddef>

namespace orc {

struct Blas {
    bool on = false;
    void* handle = nullptr;
    // cblas enums: ColMajor 102, NoTrans 111, Trans 112, Upper 121, Lower 122, NonUnit 131, Left 141
    void (*dgemv)(int order, int trans, int m, int n, double alpha, const double* A, int lda, const double* x, int incx, double beta,
                  double* y, int incy) = nullptr;
    void (*dtrsv)(int order, int uplo, int trans, int diag, int n, const double* A, int lda, double* x, int incx) = nullptr;
    void (*dtrsm)(int order, int side, int uplo, int trans, int diag, int m, int n, double alpha, const double* A, int lda, double* B,
                  int ldb) = nullptr;
    void (*dsyrk)(int order, int uplo, int trans, int n, int k, double alpha, const double* A, int lda, double beta, double* C, int ldc) = nullptr;
    void (*dpotrf)(const char* uplo, const int* n, double* A, const int* lda, int* info) = nullptr;
    void (*set_threads)(int) = nullptr;
    int threads_all = 1;
};

inline Blas& blas() {
    static Blas b;
    return b;
}

// returns 0 on success, else the 1-based index of the symbol that is missing (or -1: the library did not load)
inline int blas_load(const char* path, int threads_all) {
    Blas& b = blas();
    if (b.handle) { b.on = true; b.threads_all = threads_all; return 0; }
    void* h = dlopen(path, RTLD_NOW | RTLD_LOCAL);
    if (!h) return -1;
    const char* names[6] = {"scipy_cblas_dgemv", "scipy_cblas_dtrsv", "scipy_cblas_dtrsm", "scipy_cblas_dsyrk", "scipy_dpotrf_",
                            "scipy_openblas_set_num_threads"};
    void* p[6];
    for (int k = 0; k < 6; ++k) {
        p[k] = dlsym(h, names[k]);
        if (!p[k]) return k + 1;
    }
    b.dgemv = (decltype(b.dgemv))p[0];
    b.dtrsv = (decltype(b.dtrsv))p[1];
    b.dtrsm = (decltype(b.dtrsm))p[2];
    b.dsyrk = (decltype(b.dsyrk))p[3];
    b.dpotrf = (decltype(b.dpotrf))p[4];
    b.set_threads = (decltype(b.set_threads))p[5];
    b.handle = h;
    b.threads_all = threads_all;
    b.on = true;
    return 0;
}
inline void blas_off() { blas().on = false; }

// bracket an `omp parallel for` over items: the BLAS calls inside it must not spawn their own threads
struct BlasSerialScope {
    BlasSerialScope() { if (blas().on) blas().set_threads(1); }
    ~BlasSerialScope() { if (blas().on) blas().set_threads(blas().threads_all); }
};

}  // namespace orc
