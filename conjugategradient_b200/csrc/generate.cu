// generate.cu — synthetic stencil matrices assembled directly in device memory, one z-slab (shard) at a time.
// "on-device stencil generators, so 512^3 / 27-point never exists on the host" (the global 27-point
// 512^3 matrix has 3.6e9 non-zeros: more than int32 and 44 GB).  The reference has only host generators
// (MgcgMain.cs:51-104, SimpleConjugateGradient.cu:163-197); these produce the matrices BASELINE.json's configs name, with
// the reference's row layout: diagonal first (MgcgMain.cs:59-60), then the neighbours, 0-based GLOBAL column indices,
// row offsets rebased to the shard (Mgcg.cu:73), lexicographic idx = x + nx*(y + ny*z).
// Bit-identical to conjugategradient_b200/generators.py (poisson3d7 / poisson3d27_varcoef), which the tests check.
// Row offsets are closed-form prefix counts, so there is no scan and no second pass.
#include "common.cuh"

namespace mgcg {

// ---- 7-point: entries of the rows before local row li (x, y, zl known)
__device__ __forceinline__ long long p7_prefix(long long li, int x, int y, int zl, int nx, int ny, int nz, int z0, int z1)
{
    const long long line = li / nx, plane = li / ((long long)nx * ny);
    long long miss = 0;
    miss += line + (x > 0 ? 1 : 0);                              // rows with x == 0
    miss += line;                                                // rows with x == nx-1
    miss += plane * nx + (y == 0 ? x : nx);                      // rows with y == 0
    miss += plane * nx + (y == ny - 1 ? x : 0);                  // rows with y == ny-1
    if (z0 == 0) miss += (zl == 0) ? li : (long long)nx * ny;    // rows with z == 0
    if (z1 == nz) {                                              // rows with z == nz-1
        if (zl == z1 - z0 - 1) miss += li - (long long)zl * nx * ny;
        else if (zl >= z1 - z0) miss += (long long)nx * ny;      // (the end pointer)
    }
    return 7 * li - miss;
}

__global__ void __launch_bounds__(256) k_gen_poisson7(int nx, int ny, int nz, int z0, int z1, double *__restrict__ val,
                                                      int *__restrict__ rowptr, int *__restrict__ col)
{
    const long long rows = (long long)nx * ny * (z1 - z0);
    for (long long li = (long long)blockIdx.x * blockDim.x + threadIdx.x; li <= rows; li += (long long)gridDim.x * blockDim.x) {
        const int x = (int)(li % nx), y = (int)((li / nx) % ny), zl = (int)(li / ((long long)nx * ny));
        const long long p = p7_prefix(li, x, y, zl, nx, ny, nz, z0, z1);
        rowptr[li] = (int)p;
        if (li == rows) break;
        const int z = z0 + zl;
        const long long g = li + (long long)z0 * nx * ny;
        long long q = p;
        val[q] = 6.0; col[q] = (int)g; ++q;
        if (x > 0) { val[q] = -1.0; col[q] = (int)(g - 1); ++q; }
        if (x < nx - 1) { val[q] = -1.0; col[q] = (int)(g + 1); ++q; }
        if (y > 0) { val[q] = -1.0; col[q] = (int)(g - nx); ++q; }
        if (y < ny - 1) { val[q] = -1.0; col[q] = (int)(g + nx); ++q; }
        if (z > 0) { val[q] = -1.0; col[q] = (int)(g - (long long)nx * ny); ++q; }
        if (z < nz - 1) { val[q] = -1.0; col[q] = (int)(g + (long long)nx * ny); ++q; }
    }
}

// ---- 27-point variable coefficient: the count of a row is a product cx(x)*cy(y)*cz(z), so the prefix separates
__device__ __forceinline__ long long pre1(int x, int n) { return (long long)x + (x > 1 ? x - 1 : 0) + (x < n - 1 ? x : n - 1); }   // sum_{x'<x} c(x')
__device__ __forceinline__ int cnt1(int x, int n) { return 1 + (x > 0 ? 1 : 0) + (x < n - 1 ? 1 : 0); }

struct Gen27 {
    int nx, ny, nz, z0, z1, block, bx, by;
    double w[26];          // weight of offset o (dz-major, dy, dx order without the centre), computed on the host like generators.py
};

__device__ __forceinline__ double kappa_at(const Gen27 &g, const double *__restrict__ kblk, int x, int y, int z)
{
    x = min(max(x, 0), g.nx - 1); y = min(max(y, 0), g.ny - 1); z = min(max(z, 0), g.nz - 1);
    return kblk[(x / g.block) + g.bx * ((y / g.block) + g.by * (z / g.block))];
}

__global__ void __launch_bounds__(256) k_gen_poisson27(Gen27 g, const double *__restrict__ kblk, double *__restrict__ val,
                                                       int *__restrict__ rowptr, int *__restrict__ col)
{
    const long long rows = (long long)g.nx * g.ny * (g.z1 - g.z0);
    const long long sx = pre1(g.nx, g.nx), sy = pre1(g.ny, g.ny);
    for (long long li = (long long)blockIdx.x * blockDim.x + threadIdx.x; li <= rows; li += (long long)gridDim.x * blockDim.x) {
        const int x = (int)(li % g.nx), y = (int)((li / g.nx) % g.ny), z = g.z0 + (int)(li / ((long long)g.nx * g.ny));
        // z == z1 only for the end pointer (x = y = 0)
        const long long before_planes = (pre1(min(z, g.nz), g.nz) - pre1(g.z0, g.nz)) * sx * sy;
        long long p = before_planes;
        if (li < rows) p += (long long)cnt1(z, g.nz) * (pre1(y, g.ny) * sx + (long long)cnt1(y, g.ny) * pre1(x, g.nx));
        rowptr[li] = (int)p;
        if (li == rows) break;
        const long long gi = li + (long long)g.z0 * g.nx * g.ny;
        const double ks = kappa_at(g, kblk, x, y, z);
        double diag = 0.0;
        long long q = p + 1;
        int o = 0;
        for (int dz = -1; dz <= 1; ++dz)
            for (int dy = -1; dy <= 1; ++dy)
                for (int dx = -1; dx <= 1; ++dx) {
                    if (dx == 0 && dy == 0 && dz == 0) continue;
                    const double t = __dmul_rn(sqrt(__dmul_rn(ks, kappa_at(g, kblk, x + dx, y + dy, z + dz))), g.w[o]);
                    diag = __dadd_rn(diag, t);       // full weight incl. the eliminated Dirichlet neighbours
                    const int xx = x + dx, yy = y + dy, zz = z + dz;
                    if (xx >= 0 && xx < g.nx && yy >= 0 && yy < g.ny && zz >= 0 && zz < g.nz) {
                        val[q] = -t; col[q] = (int)(gi + dx + (long long)g.nx * (dy + (long long)g.ny * dz)); ++q;
                    }
                    ++o;
                }
        val[p] = diag; col[p] = (int)gi;
    }
}

static int gen_grid(long long rows, int sm)
{
    long long g = (rows + 256) / 256;
    if (g > (long long)sm * 32) g = (long long)sm * 32;
    return (int)(g < 1 ? 1 : g);
}

}  // namespace mgcg

using namespace mgcg;

extern "C" long long mgcg_generate_poisson3d7_nnz(int nx, int ny, int nz, int z0, int z1)
{
    if (nx < 1 || ny < 1 || nz < 1 || z0 < 0 || z1 > nz || z0 >= z1) return -1;
    const long long nzl = z1 - z0, rows = (long long)nx * ny * nzl;
    return 7 * rows - 2 * (long long)ny * nzl - 2 * (long long)nx * nzl - (long long)nx * ny * ((z0 == 0) + (z1 == nz));
}

extern "C" int mgcg_generate_poisson3d7(mgcg_ctx *c, int nx, int ny, int nz, int z0, int z1, double *d_elements, int *d_row_offsets,
                                        int *d_column_indeces)
{
    const long long nnz = mgcg_generate_poisson3d7_nnz(nx, ny, nz, z0, z1);
    if (!c || nnz < 0 || !d_elements || !d_row_offsets || !d_column_indeces) { set_error("mgcg_generate_poisson3d7: bad argument"); return MGCG_ERR_INVALID; }
    if (nnz >= (1ll << 31) || (long long)nx * ny * nz >= (1ll << 31)) { set_error("mgcg_generate_poisson3d7: shard exceeds int32 (%lld non-zeros): use more shards", nnz); return MGCG_ERR_UNSUPPORTED; }
    MGCG_CUDA(cudaSetDevice(c->device));
    const long long rows = (long long)nx * ny * (z1 - z0);
    k_gen_poisson7<<<gen_grid(rows, c->sm_count), 256, 0, c->stream>>>(nx, ny, nz, z0, z1, d_elements, d_row_offsets, d_column_indeces);
    MGCG_LAUNCH_CHECK();
    return MGCG_OK;
}

extern "C" long long mgcg_generate_poisson3d27_nnz(int nx, int ny, int nz, int z0, int z1)
{
    if (nx < 1 || ny < 1 || nz < 1 || z0 < 0 || z1 > nz || z0 >= z1) return -1;
    auto pre = [](long long x, long long n) { return x + (x > 1 ? x - 1 : 0) + (x < n - 1 ? x : n - 1); };
    return (pre(z1, nz) - pre(z0, nz)) * pre(nx, nx) * pre(ny, ny);
}

/* kblk: device array of the cell coefficients, ((nx+block-1)/block) * ((ny+block-1)/block) * ((nz+block-1)/block) entries */
extern "C" int mgcg_generate_poisson3d27(mgcg_ctx *c, int nx, int ny, int nz, int z0, int z1, int block, const double *d_kblk,
                                         const double *aniso3, double *d_elements, int *d_row_offsets, int *d_column_indeces)
{
    const long long nnz = mgcg_generate_poisson3d27_nnz(nx, ny, nz, z0, z1);
    if (!c || nnz < 0 || block < 1 || !d_kblk || !aniso3 || !d_elements || !d_row_offsets || !d_column_indeces) {
        set_error("mgcg_generate_poisson3d27: bad argument");
        return MGCG_ERR_INVALID;
    }
    if (nnz >= (1ll << 31) || (long long)nx * ny * nz >= (1ll << 31)) { set_error("mgcg_generate_poisson3d27: shard exceeds int32 (%lld non-zeros): use more shards", nnz); return MGCG_ERR_UNSUPPORTED; }
    MGCG_CUDA(cudaSetDevice(c->device));
    Gen27 g;
    g.nx = nx; g.ny = ny; g.nz = nz; g.z0 = z0; g.z1 = z1; g.block = block;
    g.bx = (nx + block - 1) / block; g.by = (ny + block - 1) / block;
    int o = 0;
    for (int dz = -1; dz <= 1; ++dz)
        for (int dy = -1; dy <= 1; ++dy)
            for (int dx = -1; dx <= 1; ++dx) {
                if (!dx && !dy && !dz) continue;
                const int d2 = dx * dx + dy * dy + dz * dz;                 // generators.py w(): 1 face, 2 edge, 3 corner
                const double a = (abs(dx) * aniso3[0] + abs(dy) * aniso3[1] + abs(dz) * aniso3[2]) / d2;
                g.w[o++] = a / d2;
            }
    const long long rows = (long long)nx * ny * (z1 - z0);
    k_gen_poisson27<<<gen_grid(rows, c->sm_count), 256, 0, c->stream>>>(g, d_kblk, d_elements, d_row_offsets, d_column_indeces);
    MGCG_LAUNCH_CHECK();
    return MGCG_OK;
}
