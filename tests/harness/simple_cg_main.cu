// simple_cg_main.cu — the caller side of SolveGpu(thrust::device_vector...) (include/mgcg.hpp), shaped like the reference's
// SimpleConjugateGradient program (SimpleConjugateGradient.cu:128-245): tridiagonal [1 2 1] system, b_i = i*i*0.5, host arrays
// copied into device vectors, one SolveGpu call, x copied back.  usage: simple_cg_main N  -> prints iteration, residual, x[0..min(N,16))
// and whether SolveGpu_c on the same device arrays returns the same bits.
#include <cstdio>
#include <cstdlib>
#include <vector>

#include <thrust/copy.h>
#include <thrust/device_vector.h>

#include "mgcg.hpp"

int main(int argc, char **argv)
{
    const int N = argc > 1 ? atoi(argv[1]) : 10;
    std::vector<double> val; std::vector<int> col, rowptr(1, 0);
    for (int i = 0; i < N; ++i) {
        val.push_back(2); col.push_back(i);                               // diagonal first, as the reference stores it
        if (i > 0) { val.push_back(1); col.push_back(i - 1); }
        if (i < N - 1) { val.push_back(1); col.push_back(i + 1); }
        rowptr.push_back((int)val.size());
    }
    std::vector<double> b(N), x(N), x2(N);
    for (int i = 0; i < N; ++i) b[i] = i * i * 0.5;                        // int32 product, then * 0.5 (SimpleConjugateGradient.cu:188)
    thrust::device_vector<double> dval(val.begin(), val.end()), db(b.begin(), b.end()), dx(N, 7.0), dx2(N, 7.0);   // x0 != 0 on purpose: SolveGpu zeroes it
    thrust::device_vector<int> dcol(col.begin(), col.end()), drp(rowptr.begin(), rowptr.end());
    int it = -1, it2 = -1; double res = -1, res2 = -1;
    SolveGpu(dval, drp, dcol, dx, db, (int)val.size(), N, 1e-8, 0, N, it, res);
    thrust::copy(dx.begin(), dx.end(), x.begin());
    SolveGpu_c(thrust::raw_pointer_cast(dval.data()), thrust::raw_pointer_cast(drp.data()), thrust::raw_pointer_cast(dcol.data()),
               thrust::raw_pointer_cast(dx2.data()), thrust::raw_pointer_cast(db.data()), (int)val.size(), N, 1e-8, 0, N, &it2, &res2);
    thrust::copy(dx2.begin(), dx2.end(), x2.begin());
    bool same = (it == it2) && (res == res2);
    for (int i = 0; i < N; ++i) same = same && (x[i] == x2[i]);
    printf("iteration=%d residual=%.17g same_as_SolveGpu_c=%d\nx=", it, res, (int)same);
    for (int i = 0; i < N && i < 16; ++i) printf("%.10f ", x[i]);
    printf("\n");
    return same ? 0 : 1;
}
