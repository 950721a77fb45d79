/*
 * mgcg_compat.h — the 32 export names of the reference's MgcgGpu.dll (extern "C" __declspec(dllexport) _stdcall;
 * on x86-64 Linux that is the default C calling convention) plus a C entry for SimpleConjugateGradient's SolveGpu.
 * Argument order and meaning are the reference's; `int&`/`double&` are pointers at ABI level (C# `out int`).
 * The handle types are opaque: what the reference filled with cuBLAS/cuSPARSE handles is a context of this library.
 */
#ifndef MGCG_COMPAT_H
#define MGCG_COMPAT_H
#ifdef __cplusplus
extern "C" {
#endif

typedef struct MgcgBlas MgcgBlas;         /* was cublasHandle_t*      (Runtime.cu:20) */
typedef struct MgcgSparse MgcgSparse;     /* was cusparseHandle_t*    (Runtime.cu:34) */
typedef struct MgcgMatDescr MgcgMatDescr; /* was cusparseMatDescr_t*  (Runtime.cu:48) */
typedef struct MgcgVector MgcgVector;       /* was thrust::device_vector<double>* (Vector_Double.hpp:5) */
typedef struct MgcgVectorInt MgcgVectorInt; /* was thrust::device_vector<int>*    (Vector_Int.hpp) */

/* ---- Runtime.cu:7-62 */
int           GetDeviceCount(void);
void          SetDevice(int deviceID);
MgcgBlas     *CreateBlas(void);
void          DestroyBlas(MgcgBlas *cublas);
MgcgSparse   *CreateSparse(void);
void          DestroySparse(MgcgSparse *cusparse);
MgcgMatDescr *CreateMatDescr(void);
void          DestroyMatDescr(MgcgMatDescr *matDescr);

/* ---- Vector_Double.cu:7-45 */
MgcgVector *Create_Double(int size);
void        CopyToArray_Double(const MgcgVector *source, double destination[], int count, int sourceOffset, int destinationOffset);
void        CopyFromArray_Double(MgcgVector *destination, const double source[], int count, int sourceOffset, int destinationOffset);
void        Delete_Double(MgcgVector *vec);
double     *ToRawPtr_Double(MgcgVector *vec);
/* Vector_Double.cu:41-45 copies `count` BYTES (bug); this copies `count` elements. */
void        CopyFromDevice_Double(const double *source, double *destination, int count, int sourceOffset, int destinationOffset);

/* ---- Vector_Int.cu:9-40 */
MgcgVectorInt *Create_Int(int size);
void           CopyToArray_Int(const MgcgVectorInt *source, int destination[], int count, int sourceOffset, int destinationOffset);
void           CopyFromArray_Int(MgcgVectorInt *destination, int source[], int count, int sourceOffset, int destinationOffset);
void           Delete_Int(MgcgVectorInt *vec);
int           *ToRawPtr_Int(MgcgVectorInt *vec);

/* ---- Mgcg.cu:10-54 (raw device pointers) */
void   CsrMV(MgcgSparse *cusparse, MgcgMatDescr *matDescr, double *y,
             const double *elements, const int *rowOffsets, const int *columnIndeces, const double *x,
             int elementsCount, int rowCount, int columnCount, double alpha, double beta);
void   Axpy(MgcgBlas *cublas, double *y, const double *x, int count, double alpha);
double Dot(MgcgBlas *cublas, double *y, const double *x, int count);
void   Scal(MgcgBlas *cublas, double *x, double alpha, int count);
void   Copy(MgcgBlas *cublas, double *y, const double *x, int count, int yOffset, int xOffset);

/* ---- Mgcg.cu:57-113 multi-GPU upload and host-staged halo */
void Initialize(const double elements[], const int rowOffsets[], const int columnIndeces[],
                const double x[], const double b[],
                MgcgVector *elementsVector, MgcgVectorInt *rowOffsetsVector, MgcgVectorInt *columnIndecesVector,
                MgcgVector *xVector, MgcgVector *bVector, MgcgVector *pVector,
                int *minJ, int *maxJ, int count,
                int countForDevice, int offsetForDevice, int elementCountForDevice, int elementOffsetForDevice);
void P2Host(MgcgVector *pVector, double p[], int thisCount, int thisOffset, int lastCount, int nextCount);
void P2Device(MgcgVector *pVector, double p[], int thisCount, int thisOffset, int lastCount, int nextCount);

/* ---- Mgcg.cu:116-198 split-phase CG (each returns this device's partial dot) */
double Solve0(MgcgBlas *cublas, MgcgSparse *cusparse, MgcgMatDescr *matDescr,
              MgcgVector *elementsVector, MgcgVectorInt *rowOffsetsVector, MgcgVectorInt *columnIndecesVector,
              MgcgVector *xVector, MgcgVector *bVector, MgcgVector *ApVector, MgcgVector *pVector, MgcgVector *rVector,
              int count, int countForDevice, int offsetForDevice, int elementsCountForDevice);
double Solve1(MgcgBlas *cublas, MgcgSparse *cusparse, MgcgMatDescr *matDescr,
              MgcgVector *elementsVector, MgcgVectorInt *rowOffsetsVector, MgcgVectorInt *columnIndecesVector,
              MgcgVector *ApVector, MgcgVector *pVector,
              int count, int countForDevice, int offsetForDevice, int elementsCountForDevice);
double Solve2(MgcgBlas *cublas, double alpha, MgcgVector *xVector, MgcgVector *ApVector, MgcgVector *pVector,
              MgcgVector *rVector, int countForDevice, int offsetForDevice);
void   Solve3(MgcgBlas *cublas, double beta, MgcgVector *pVector, MgcgVector *rVector, int countForDevice, int offsetForDevice);

/* ---- Mgcg.cu:201-270 whole single-GPU CG; x = initial guess; *iteration = iterations performed */
void Solve(MgcgBlas *cublas, MgcgSparse *cusparse, MgcgMatDescr *matDescr,
           MgcgVector *elementsVector, MgcgVectorInt *rowOffsetsVector, MgcgVectorInt *columnIndecesVector,
           MgcgVector *xVector, MgcgVector *bVector, MgcgVector *ApVector, MgcgVector *pVector, MgcgVector *rVector,
           int elementsCount, int count, double allowableResidual, int minIteration, int maxIteration,
           int *iteration, double *residual);

/* ---- SimpleConjugateGradient.cu:12-126 SolveGpu on raw device pointers (the C++ overload taking
 *      thrust::device_vector is the header-only include/mgcg.hpp, tests/test_solvegpu_overload.py): x0 forced to 0 (:53),
 *      strict min-iteration (:107). */
void SolveGpu_c(const double *elements, const int *rowOffsets, const int *columnIndeces,
                double *x, const double *b, int elementsCount, int count,
                double allowableResidual, int minIteration, int maxIteration, int *iteration, double *residual);

/* Extension: make the compat `Solve`/`SolveGpu_c` use the multigrid preconditioner (0 = reference-parity plain CG).
 * nx,ny,nz = optional grid hint (0 = none).  Process-wide; default off. */
void MgcgCompatSetPreconditioner(int precond, int nx, int ny, int nz);

#ifdef __cplusplus
}
#endif
#endif
