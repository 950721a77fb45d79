/*
 * mgcg.hpp — the C++ face of SolveGpu: the overload on thrust::device_vector that the reference's SimpleConjugateGradient
 * program defines and calls (SimpleConjugateGradient/SimpleConjugateGradient/SimpleConjugateGradient.cu:12-22, call site :231).
 * Same argument order, same value / reference passing (the matrix and b by value, x by reference, iteration and residual as
 * output references), same behaviour: x is zeroed first (:53), the minimum-iteration test is strict (:107), `iteration` is
 * the number of iterations done and `residual` = sqrt(r.r) at exit (:106,:119-120).
 * Header-only: unwraps the vectors and calls SolveGpu_c of libMgcgGpu.so (mgcg_compat.h).  Needs nvcc (thrust's CUDA backend);
 * nothing of thrust crosses the library boundary.
 */
#ifndef MGCG_HPP
#define MGCG_HPP

#include <thrust/device_vector.h>

#include "mgcg_compat.h"

inline void SolveGpu(
    thrust::device_vector<double> elementsVector, thrust::device_vector<int> rowOffsetsVector, thrust::device_vector<int> columnIndecesVector,
    thrust::device_vector<double> &xVector, thrust::device_vector<double> bVector,
    const int elementsCount, const int count,
    const double allowableResidual, const int minIteration, const int maxIteration,
    int &iteration, double &residual)
{
    SolveGpu_c(thrust::raw_pointer_cast(elementsVector.data()), thrust::raw_pointer_cast(rowOffsetsVector.data()),
               thrust::raw_pointer_cast(columnIndecesVector.data()), thrust::raw_pointer_cast(xVector.data()),
               thrust::raw_pointer_cast(bVector.data()), elementsCount, count, allowableResidual, minIteration, maxIteration,
               &iteration, &residual);
}

#endif /* MGCG_HPP */
