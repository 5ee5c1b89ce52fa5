// ORACLE - TEST INFRASTRUCTURE ONLY.
// Thin C entry points around the reference's own LAPACK-free LU, compiled from the sources where
// they lie (external/anyode/include/anyode/anyode_blasless.hpp:10-95 under /root/reference) into
// oracle/_ref/libanyode_blasless.so by oracle/build.py.  No reference source is copied.
#include <anyode/anyode_blasless.hpp>
#include <cstdint>

extern "C" int ref_anyode_lu_solve(int64_t nsys, int n, const double *A_rowmajor, const double *b, double *x,
                                   int32_t *info_out) {
    AnyODE::getrf_callback<double> getrf;
    AnyODE::getrs_callback<double> getrs;
    double *a = new double[n * n];
    int *ipiv = new int[n];
    for (int64_t s = 0; s < nsys; ++s) {
        for (int r = 0; r < n; ++r)
            for (int c = 0; c < n; ++c) a[c * n + r] = A_rowmajor[s * n * n + r * n + c];  // column major
        int info = 0, lda = n, nrhs = 1;
        getrf(&n, &n, a, &lda, ipiv, &info);
        for (int i = 0; i < n; ++i) x[s * n + i] = b[s * n + i];
        int info2 = 0;
        char trans = 'N';
        getrs(&trans, &n, &nrhs, a, &lda, ipiv, x + s * n, &lda, &info2);
        info_out[s] = info;
    }
    delete[] a;
    delete[] ipiv;
    return 0;
}
