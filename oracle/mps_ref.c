/*
 * oracle/mps_ref.c — TEST INFRASTRUCTURE ONLY (never linked into, imported by or called from the product path).
 *
 * CPU statement of the particle (MPS) pressure-Poisson assembly that SURVEY.md §8(f)2 asks to move onto the device:
 * neighbour search by cell list -> CSR.  The reference builds this matrix in its (absent) MPS caller and hands the solver a
 * CSR with at most 160 entries per row (Mgcg/cuBlas/Mgcg/MgcgMain.cs:20); there is no reference
 * code for the search itself, so this file DEFINES the semantics a device kernel has to reproduce bit for bit:
 *
 *   cells      : cubes of edge `re` on the bounding box; cell id = cx + ncx*(cy + ncy*cz); particles are numbered in
 *                cell-id order (stable: ties keep the input order) — ref_mps_cell_order.
 *   neighbours : j != i with r_ij < re, r = sqrt((dx*dx + dy*dy) + dz*dz) (no contraction; build with -ffp-contract=off).
 *                Scanning the 27 adjacent cells by (dz, dy) ascending and the three x-adjacent cells as one contiguous
 *                run of particle numbers yields the row already in ascending column order: no per-row sort.
 *   values     : a_ij = -(re / r - 1); a_ii = (sum of the row's weights, added in ascending j) + shift_abs.
 *   layout     : the reference's — diagonal first, then the neighbours (MgcgMain.cs:59-60), 0-based columns, int32.
 * "parity unpinned": no reference implementation, no fixture; cross-checked against a kd-tree search in tests/test_oracle_mps.py.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

typedef struct { double x0, y0, z0; int ncx, ncy, ncz; } mps_grid;

static void mps_make_grid(int n, const double *pos, double re, mps_grid *g)
{
    double lo[3] = {pos[0], pos[1], pos[2]}, hi[3] = {pos[0], pos[1], pos[2]};
    for (int i = 1; i < n; i++)
        for (int d = 0; d < 3; d++) {
            double v = pos[3 * (size_t)i + d];
            if (v < lo[d]) lo[d] = v;
            if (v > hi[d]) hi[d] = v;
        }
    g->x0 = lo[0]; g->y0 = lo[1]; g->z0 = lo[2];
    g->ncx = (int)floor((hi[0] - lo[0]) / re) + 1;
    g->ncy = (int)floor((hi[1] - lo[1]) / re) + 1;
    g->ncz = (int)floor((hi[2] - lo[2]) / re) + 1;
}

static int mps_cell_of(const mps_grid *g, const double *p, double re)
{
    int cx = (int)floor((p[0] - g->x0) / re), cy = (int)floor((p[1] - g->y0) / re), cz = (int)floor((p[2] - g->z0) / re);
    return cx + g->ncx * (cy + g->ncy * cz);
}

/* order[k] = input index of the particle that gets number k (counting sort by cell id, stable).
 * cell_start has ncells + 1 entries (caller passes a buffer of that size, or NULL); returns ncells, < 0 on error. */
int ref_mps_cell_order(int n, const double *pos, double re, int *order, int *cell_start_out, int cell_start_cap)
{
    if (n <= 0 || re <= 0) return -1;
    mps_grid g;
    mps_make_grid(n, pos, re, &g);
    long long nc = (long long)g.ncx * g.ncy * g.ncz;
    if (nc <= 0 || nc >= (1ll << 31) - 1) return -1;
    int *start = (int *)calloc((size_t)nc + 1, sizeof(int));
    int *cell = (int *)malloc(sizeof(int) * (size_t)n);
    if (!start || !cell) { free(start); free(cell); return -1; }
    for (int i = 0; i < n; i++) { cell[i] = mps_cell_of(&g, pos + 3 * (size_t)i, re); start[cell[i] + 1]++; }
    for (long long c = 0; c < nc; c++) start[c + 1] += start[c];
    int *fill = (int *)malloc(sizeof(int) * (size_t)nc);
    if (!fill) { free(start); free(cell); return -1; }
    memcpy(fill, start, sizeof(int) * (size_t)nc);
    for (int i = 0; i < n; i++) order[fill[cell[i]]++] = i;
    if (cell_start_out && cell_start_cap >= nc + 1) memcpy(cell_start_out, start, sizeof(int) * ((size_t)nc + 1));
    free(fill); free(cell); free(start);
    return (int)nc;
}

/* pos MUST already be numbered in cell order (apply ref_mps_cell_order first).  Two passes like a device build:
 * pass 1 (elements == NULL): row_offsets only; returns the entry count.  pass 2: fills elements / column_indeces.
 * Returns < 0 on error (-2: a row exceeds max_per_row, the reference's container limit, SparseMatrix.cs:44-49). */
long long ref_mps_assemble(int n, const double *pos, double re, double shift_abs, int max_per_row,
                           int *row_offsets, double *elements, int *column_indeces)
{
    if (n <= 0 || re <= 0) return -1;
    mps_grid g;
    mps_make_grid(n, pos, re, &g);
    long long nc = (long long)g.ncx * g.ncy * g.ncz;
    int *start = (int *)calloc((size_t)nc + 1, sizeof(int));
    if (!start) return -1;
    int prev = -1;
    for (int i = 0; i < n; i++) {
        int c = mps_cell_of(&g, pos + 3 * (size_t)i, re);
        if (c < prev) { free(start); return -3; }                 /* not in cell order */
        prev = c;
        start[c + 1]++;
    }
    for (long long c = 0; c < nc; c++) start[c + 1] += start[c];
    long long o = 0;
    for (int i = 0; i < n; i++) {
        const double *pi = pos + 3 * (size_t)i;
        int cx = (int)floor((pi[0] - g.x0) / re), cy = (int)floor((pi[1] - g.y0) / re), cz = (int)floor((pi[2] - g.z0) / re);
        long long row0 = o;
        double diag = 0;
        row_offsets[i] = (int)o;
        o++;                                                      /* the diagonal's slot */
        for (int dz = -1; dz <= 1; dz++) {
            int zz = cz + dz;
            if (zz < 0 || zz >= g.ncz) continue;
            for (int dy = -1; dy <= 1; dy++) {
                int yy = cy + dy;
                if (yy < 0 || yy >= g.ncy) continue;
                int xa = cx > 0 ? cx - 1 : 0, xb = cx < g.ncx - 1 ? cx + 1 : g.ncx - 1;
                long long base = (long long)g.ncx * (yy + (long long)g.ncy * zz);
                int j0 = start[base + xa], j1 = start[base + xb + 1];      /* one contiguous run of particle numbers */
                for (int j = j0; j < j1; j++) {
                    if (j == i) continue;
                    const double *pj = pos + 3 * (size_t)j;
                    double dx = pi[0] - pj[0], dyv = pi[1] - pj[1], dzv = pi[2] - pj[2];
                    double r = sqrt((dx * dx + dyv * dyv) + dzv * dzv);
                    if (!(r < re)) continue;
                    double w = re / r - 1.0;
                    diag += w;
                    if (elements) { elements[o] = -w; column_indeces[o] = j; }
                    o++;
                }
            }
        }
        if (o - row0 > max_per_row) { free(start); return -2; }
        if (elements) { elements[row0] = diag + shift_abs; column_indeces[row0] = i; }
        if (o >= (1ll << 31)) { free(start); return -1; }
    }
    row_offsets[n] = (int)o;
    free(start);
    return o;
}
