// mgcg_main.cpp — the caller side of the drop-in boundary in C++: the flow of the reference's harness
// program MgcgMain.Main (Mgcg/cuBlas/Mgcg/MgcgMain.cs:41-178) and of its three solver classes
//   ConjugateGradientCpu          (ConjugateGradientCpu.cs:45-98)    -> the CPU oracle (oracle/cg_ref.c: test infrastructure)
//   ConjugateGradientSingleGpu    (ConjugateGradientSingleGpu.cs:98-178)
//   ConjugateGradientParallelGpu  (ConjugateGradientParallelGpu.cs:264-594), one host thread per device like its Parallel.For
// driving the UNCHANGED export names of MgcgGpu.dll (include/mgcg_compat.h) in libMgcgGpu.so: generate -> Initialize ->
// timed Solve -> Read -> CPU-vs-GPU comparison with the harness's own 1 % rule (MgcgMain.cs:135).
// This file is test infrastructure (it links the oracle); it is what a maintainer's C# host does through P/Invoke.
//
//   mgcg_main <libMgcgGpu.so> <liboracle.so> [count] [max_nonzero] [min_iteration] [devices]
// Prints one line per solver and "HARNESS_OK" when every comparison passes; exit code 0/1.
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <dlfcn.h>
#include <functional>
#include <thread>
#include <vector>

#include "../../include/mgcg_compat.h"

#define LOAD(lib, name) name##_t name##_f = (name##_t)dlsym(lib, #name); if (!name##_f) { fprintf(stderr, "missing symbol %s\n", #name); return 1; }

typedef int (*GetDeviceCount_t)(void);
typedef void (*SetDevice_t)(int);
typedef MgcgBlas *(*CreateBlas_t)(void);
typedef void (*DestroyBlas_t)(MgcgBlas *);
typedef MgcgSparse *(*CreateSparse_t)(void);
typedef void (*DestroySparse_t)(MgcgSparse *);
typedef MgcgMatDescr *(*CreateMatDescr_t)(void);
typedef void (*DestroyMatDescr_t)(MgcgMatDescr *);
typedef MgcgVector *(*Create_Double_t)(int);
typedef void (*CopyToArray_Double_t)(const MgcgVector *, double *, int, int, int);
typedef void (*CopyFromArray_Double_t)(MgcgVector *, const double *, int, int, int);
typedef void (*Delete_Double_t)(MgcgVector *);
typedef MgcgVectorInt *(*Create_Int_t)(int);
typedef void (*CopyFromArray_Int_t)(MgcgVectorInt *, int *, int, int, int);
typedef void (*Delete_Int_t)(MgcgVectorInt *);
typedef void (*Initialize_t)(const double *, const int *, const int *, const double *, const double *, MgcgVector *, MgcgVectorInt *,
                             MgcgVectorInt *, MgcgVector *, MgcgVector *, MgcgVector *, int *, int *, int, int, int, int, int);
typedef void (*P2Host_t)(MgcgVector *, double *, int, int, int, int);
typedef void (*P2Device_t)(MgcgVector *, double *, int, int, int, int);
typedef double (*Solve0_t)(MgcgBlas *, MgcgSparse *, MgcgMatDescr *, MgcgVector *, MgcgVectorInt *, MgcgVectorInt *, MgcgVector *, MgcgVector *,
                           MgcgVector *, MgcgVector *, MgcgVector *, int, int, int, int);
typedef double (*Solve1_t)(MgcgBlas *, MgcgSparse *, MgcgMatDescr *, MgcgVector *, MgcgVectorInt *, MgcgVectorInt *, MgcgVector *, MgcgVector *,
                           int, int, int, int);
typedef double (*Solve2_t)(MgcgBlas *, double, MgcgVector *, MgcgVector *, MgcgVector *, MgcgVector *, int, int);
typedef void (*Solve3_t)(MgcgBlas *, double, MgcgVector *, MgcgVector *, int, int);
typedef void (*Solve_t)(MgcgBlas *, MgcgSparse *, MgcgMatDescr *, MgcgVector *, MgcgVectorInt *, MgcgVectorInt *, MgcgVector *, MgcgVector *,
                        MgcgVector *, MgcgVector *, MgcgVector *, int, int, double, int, int, int *, double *);
// oracle (CPU restatement of the C# CPU path)
typedef int (*ref_cg_solve_t)(int, const double *, const int *, const int *, double *, const double *, double *, double *, double *, double, int,
                              int, int, int, int *, double *);
typedef int (*ref_gen_banded_sin_t)(int, int, double *, int *, int *, double *, double *);

// Parallel.For(0, n, body) with one host thread per index (ConjugateGradientParallelGpu.cs:301, 436, ...)
static bool g_serial = false;
static void parallel_for(int n, const std::function<void(int)> &body)
{
    if (g_serial) { for (int i = 0; i < n; ++i) body(i); return; }
    std::vector<std::thread> th;
    for (int i = 0; i < n; ++i) th.emplace_back(body, i);
    for (auto &t : th) t.join();
}

static double now_ms()
{
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// ConjugateGradient.IsConverged (ConjugateGradient.cs:56-79): 0 continue, 1 converged, -1 throw
static int is_converged(int iteration, int min_it, int max_it, double residual, double allowable)
{
    if (iteration < min_it) return 0;
    if (iteration > max_it) return -1;
    return residual < allowable ? 1 : 0;
}

int main(int argc, char **argv)
{
    if (argc < 3) { fprintf(stderr, "usage: %s libMgcgGpu.so liboracle.so [count] [max_nonzero] [min_iteration] [devices]\n", argv[0]); return 2; }
    const int COUNT = argc > 3 ? atoi(argv[3]) : 34567 * 6;          // MgcgMain.cs:10
    const int MAX_NONZERO_COUNT = argc > 4 ? atoi(argv[4]) : 160;    // :20
    const int MIN_ITERATION = argc > 5 ? atoi(argv[5]) : 200;        // :25
    const int MAX_ITERATION = COUNT;                                 // :30
    const double ALLOWABLE_RESIDUAL = 1e-8;                          // :35
    void *gpu = dlopen(argv[1], RTLD_NOW | RTLD_LOCAL);
    if (!gpu) { fprintf(stderr, "cannot load %s: %s\n", argv[1], dlerror()); return 1; }
    void *cpu = dlopen(argv[2], RTLD_NOW | RTLD_LOCAL);
    if (!cpu) { fprintf(stderr, "cannot load %s: %s\n", argv[2], dlerror()); return 1; }
    LOAD(gpu, GetDeviceCount) LOAD(gpu, SetDevice) LOAD(gpu, CreateBlas) LOAD(gpu, DestroyBlas) LOAD(gpu, CreateSparse) LOAD(gpu, DestroySparse)
    LOAD(gpu, CreateMatDescr) LOAD(gpu, DestroyMatDescr) LOAD(gpu, Create_Double) LOAD(gpu, CopyToArray_Double) LOAD(gpu, CopyFromArray_Double)
    LOAD(gpu, Delete_Double) LOAD(gpu, Create_Int) LOAD(gpu, CopyFromArray_Int) LOAD(gpu, Delete_Int) LOAD(gpu, Initialize) LOAD(gpu, P2Host)
    LOAD(gpu, P2Device) LOAD(gpu, Solve0) LOAD(gpu, Solve1) LOAD(gpu, Solve2) LOAD(gpu, Solve3) LOAD(gpu, Solve)
    LOAD(cpu, ref_cg_solve) LOAD(cpu, ref_gen_banded_sin)
    const int physical = GetDeviceCount_f();
    int deviceCount = physical;
    if (argc > 6 && atoi(argv[6]) > 0) deviceCount = atoi(argv[6]);
    // more logical "devices" than GPUs (tests on one GPU): block d runs on GPU d % physical, and the blocks are walked by
    // one thread, because the per-device handles of the library are not meant to be shared by concurrent threads
    auto setDevice = [&](int d) { SetDevice_f(d % physical); };
    g_serial = deviceCount > physical;
    printf("N=%d devices=%d\n", COUNT, deviceCount);

    // ---- the system (MgcgMain.cs:51-104): SparseMatrix pre-sized rows x maxNonzeroPerRow (SparseMatrix.cs:33-34)
    std::vector<double> elements((size_t)COUNT * MAX_NONZERO_COUNT), b(COUNT), x0(COUNT);
    std::vector<int> columnIndeces((size_t)COUNT * MAX_NONZERO_COUNT), rowOffsets(COUNT + 1);
    const int nnz = ref_gen_banded_sin_f(COUNT, MAX_NONZERO_COUNT, elements.data(), columnIndeces.data(), rowOffsets.data(), b.data(), x0.data());

    // ---- ConjugateGradientCpu (the comparison baseline, MgcgMain.cs:113-119)
    std::vector<double> xCpu(x0), r(COUNT), p(COUNT), Ap(COUNT);
    int itCpu = 0; double resCpu = 0;
    double t0 = now_ms();
    int st = ref_cg_solve_f(COUNT, elements.data(), rowOffsets.data(), columnIndeces.data(), xCpu.data(), b.data(), r.data(), p.data(), Ap.data(),
                            ALLOWABLE_RESIDUAL, MIN_ITERATION, MAX_ITERATION, 0, 0, &itCpu, &resCpu);
    double tCpu = now_ms() - t0;
    printf("CPU          : %10.3f ms  Iteration=%d Residual=%.3e status=%d\n", tCpu, itCpu, resCpu, st);
    if (st != 0) return 1;

    // ---- ConjugateGradientSingleGpu: ctor (cs:98-119), Initialize (:133-143), Solve (:148-170), Read (:175-178)
    std::vector<double> xSingle(x0);
    int itSingle = 0; double resSingle = 0, tSingle = 0;
    {
        MgcgBlas *blas = CreateBlas_f(); MgcgSparse *sparse = CreateSparse_f(); MgcgMatDescr *descr = CreateMatDescr_f();
        MgcgVector *vE = Create_Double_f(COUNT * MAX_NONZERO_COUNT), *vX = Create_Double_f(COUNT), *vB = Create_Double_f(COUNT);
        MgcgVector *vAp = Create_Double_f(COUNT), *vP = Create_Double_f(COUNT), *vR = Create_Double_f(COUNT);
        MgcgVectorInt *vC = Create_Int_f(COUNT * MAX_NONZERO_COUNT), *vRo = Create_Int_f(COUNT + 1);
        CopyFromArray_Double_f(vE, elements.data(), nnz, 0, 0);
        CopyFromArray_Int_f(vC, columnIndeces.data(), nnz, 0, 0);
        CopyFromArray_Int_f(vRo, rowOffsets.data(), COUNT + 1, 0, 0);
        CopyFromArray_Double_f(vX, xSingle.data(), COUNT, 0, 0);
        CopyFromArray_Double_f(vB, b.data(), COUNT, 0, 0);
        int iteration = 0;
        t0 = now_ms();
        Solve_f(blas, sparse, descr, vE, vRo, vC, vX, vB, vAp, vP, vR, nnz, COUNT, ALLOWABLE_RESIDUAL, MIN_ITERATION, MAX_ITERATION, &iteration, &resSingle);
        tSingle = now_ms() - t0;
        itSingle = iteration - 1;                                   // ConjugateGradientSingleGpu.cs:168
        CopyToArray_Double_f(vX, xSingle.data(), COUNT, 0, 0);
        Delete_Double_f(vE); Delete_Double_f(vX); Delete_Double_f(vB); Delete_Double_f(vAp); Delete_Double_f(vP); Delete_Double_f(vR);
        Delete_Int_f(vC); Delete_Int_f(vRo);
        DestroyBlas_f(blas); DestroySparse_f(sparse); DestroyMatDescr_f(descr);
    }
    printf("GPU single   : %10.3f ms  Iteration=%d Residual=%.3e\n", tSingle, itSingle, resSingle);

    // ---- ConjugateGradientParallelGpu: ctor (cs:264-324), Initialize (:358-379), SyncP (:384-419), Solve (:424-565), Read (:570-585)
    const int D = deviceCount;
    std::vector<int> offsetsForDevice(D + 1, 0), minJ(D), maxJ(D);
    for (int i = 1; i < D; ++i) offsetsForDevice[i] = offsetsForDevice[i - 1] + (int)std::floor((double)COUNT / D);   // :271-277
    offsetsForDevice[D] = COUNT;
    auto countForDevice = [&](int d) { return offsetsForDevice[d + 1] - offsetsForDevice[d]; };                        // :590-594
    std::vector<double> resultsDot(D), bufferHost(COUNT), xPar(x0);
    std::vector<MgcgBlas *> blas(D); std::vector<MgcgSparse *> sparse(D); std::vector<MgcgMatDescr *> descr(D);
    std::vector<MgcgVector *> vE(D), vX(D), vB(D), vAp(D), vP(D), vR(D);
    std::vector<MgcgVectorInt *> vC(D), vRo(D);
    parallel_for(D, [&](int d) {
        setDevice(d);
        const int cnt = countForDevice(d);
        blas[d] = CreateBlas_f(); sparse[d] = CreateSparse_f(); descr[d] = CreateMatDescr_f();
        vE[d] = Create_Double_f(cnt * MAX_NONZERO_COUNT); vC[d] = Create_Int_f(cnt * MAX_NONZERO_COUNT); vRo[d] = Create_Int_f(cnt + 1);
        vX[d] = Create_Double_f(cnt); vB[d] = Create_Double_f(cnt); vAp[d] = Create_Double_f(cnt);
        vP[d] = Create_Double_f(COUNT);                                                                                // :321 full length
        vR[d] = Create_Double_f(cnt);
    });
    parallel_for(D, [&](int d) {
        setDevice(d);
        const int off = offsetsForDevice[d];
        Initialize_f(elements.data(), rowOffsets.data(), columnIndeces.data(), xPar.data(), b.data(), vE[d], vRo[d], vC[d], vX[d], vB[d], vP[d],
                     &minJ[d], &maxJ[d], COUNT, countForDevice(d), off, rowOffsets[offsetsForDevice[d + 1]] - rowOffsets[off], rowOffsets[off]);
    });
    auto SyncP = [&]() {
        for (int phase = 0; phase < 2; ++phase)
            parallel_for(D, [&](int d) {
                setDevice(d);
                const int cnt = countForDevice(d), off = offsetsForDevice[d];
                const int lastCount = d > 0 ? off - minJ[d] : 0;                                                       // :397
                const int nextCount = d < D - 1 ? maxJ[d] - cnt - off + 1 : 0;                                         // :398
                (phase == 0 ? P2Host_f : P2Device_f)(vP[d], bufferHost.data(), cnt, off, lastCount, nextCount);
            });
    };
    auto sum = [&]() { double s = 0; for (int d = 0; d < D; ++d) s += resultsDot[d]; return s; };                      // resultsDot.Sum(), device order
    auto elemCount = [&](int d) { return rowOffsets[offsetsForDevice[d + 1]] - rowOffsets[offsetsForDevice[d]]; };
    int itPar = 0; double resPar = 0;
    t0 = now_ms();
    SyncP();
    parallel_for(D, [&](int d) {
        setDevice(d);
        resultsDot[d] = Solve0_f(blas[d], sparse[d], descr[d], vE[d], vRo[d], vC[d], vX[d], vB[d], vAp[d], vP[d], vR[d], COUNT, countForDevice(d),
                                 offsetsForDevice[d], elemCount(d));
    });
    double rr = sum();
    for (itPar = 0;; ++itPar) {
        SyncP();
        parallel_for(D, [&](int d) {
            setDevice(d);
            resultsDot[d] = Solve1_f(blas[d], sparse[d], descr[d], vE[d], vRo[d], vC[d], vAp[d], vP[d], COUNT, countForDevice(d), offsetsForDevice[d], elemCount(d));
        });
        const double alpha = rr / sum();
        parallel_for(D, [&](int d) {
            setDevice(d);
            resultsDot[d] = Solve2_f(blas[d], alpha, vX[d], vAp[d], vP[d], vR[d], countForDevice(d), offsetsForDevice[d]);
        });
        const double rrNew = sum();
        resPar = std::sqrt(rrNew);
        const int c = is_converged(itPar, MIN_ITERATION, MAX_ITERATION, resPar, ALLOWABLE_RESIDUAL);
        if (c < 0) { fprintf(stderr, "GPU parallel: Iteration > MaxIteration\n"); return 1; }
        if (c) break;
        const double beta = rrNew / rr;
        parallel_for(D, [&](int d) { setDevice(d); Solve3_f(blas[d], beta, vP[d], vR[d], countForDevice(d), offsetsForDevice[d]); });
        rr = rrNew;
    }
    const double tPar = now_ms() - t0;
    parallel_for(D, [&](int d) { setDevice(d); CopyToArray_Double_f(vX[d], xPar.data(), countForDevice(d), 0, offsetsForDevice[d]); });   // Read, :570-585
    parallel_for(D, [&](int d) {
        setDevice(d);
        Delete_Double_f(vE[d]); Delete_Int_f(vC[d]); Delete_Int_f(vRo[d]); Delete_Double_f(vX[d]); Delete_Double_f(vB[d]);
        Delete_Double_f(vAp[d]); Delete_Double_f(vP[d]); Delete_Double_f(vR[d]);
        DestroyBlas_f(blas[d]); DestroySparse_f(sparse[d]); DestroyMatDescr_f(descr[d]);
    });
    printf("GPU parallel : %10.3f ms  Iteration=%d Residual=%.3e\n", tPar, itPar, resPar);

    // ---- comparison (MgcgMain.cs:127-167): |x_cpu - x_gpu| / |x_cpu| <= 1 % per element; and this repo's tighter bar
    int bad = 0;
    double worstSingle = 0, worstPar = 0, xmax = 0;
    for (int i = 0; i < COUNT; ++i) xmax = std::fmax(xmax, std::fabs(xCpu[i]));
    for (int i = 0; i < COUNT; ++i) {
        const double es = std::fabs(xCpu[i] - xSingle[i]), ep = std::fabs(xCpu[i] - xPar[i]);
        if (es / std::fabs(xCpu[i]) > 0.01 || ep / std::fabs(xCpu[i]) > 0.01) ++bad;      // the reference's own acceptance test
        worstSingle = std::fmax(worstSingle, es / xmax); worstPar = std::fmax(worstPar, ep / xmax);
    }
    printf("compare      : reference 1%% rule violations=%d  max|dx|/max|x| single=%.3e parallel=%.3e  iterations cpu/single/parallel=%d/%d/%d\n",
           bad, worstSingle, worstPar, itCpu, itSingle, itPar);
    const bool ok = bad == 0 && worstSingle <= 1e-10 && worstPar <= 1e-10 && std::abs(itSingle - itCpu) <= 1 && std::abs(itPar - itCpu) <= 1;
    printf(ok ? "HARNESS_OK\n" : "HARNESS_FAILED\n");
    return ok ? 0 : 1;
}
