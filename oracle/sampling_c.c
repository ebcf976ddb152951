/* ORACLE — TEST INFRASTRUCTURE ONLY.  Plain-C restatement of oracle/sampling.py (farthest-point
 * sampling and voxel-grid downsampling) plus the reference's per-cloud normalisation
 * (PCA/pointnet/PointNet.py:691-706), used as the timed CPU baseline and to cross-check the
 * NumPy oracle.  PARITY UNPINNED for FPS / voxel: the reference has no implementation of either
 * (SURVEY.md §0.3).  Compile with -ffp-contract=off so no FMA is formed (oracle/Makefile).
 * Never linked into the product library.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* oracle/sampling.py:fps — d = min(d, (dx*dx + dy*dy) + dz*dz), argmax lowest index on ties. */
void oracle_fps_f32(const float *pts, int64_t n, int s, int start, int32_t *out, float *dist_scratch)
{
    float *d = dist_scratch;
    for (int64_t i = 0; i < n; ++i) d[i] = INFINITY;
    int64_t cur = start;
    for (int it = 0; it < s; ++it) {
        out[it] = (int32_t)cur;
        const float cx = pts[3 * cur], cy = pts[3 * cur + 1], cz = pts[3 * cur + 2];
        float best = -1.0f;
        int64_t besti = 0;
        for (int64_t i = 0; i < n; ++i) {
            const float dx = pts[3 * i] - cx, dy = pts[3 * i + 1] - cy, dz = pts[3 * i + 2] - cz;
            float dd = dx * dx + dy * dy;
            dd = dd + dz * dz;
            float m = d[i] < dd ? d[i] : dd;
            d[i] = m;
            if (m > best) { best = m; besti = i; }
        }
        cur = besti;
    }
}

void oracle_fps_batch_f32(const float *pts, int b, int64_t n, int s, int start, int32_t *out)
{
    float *scratch = (float *)malloc(sizeof(float) * (size_t)n);
    for (int c = 0; c < b; ++c) oracle_fps_f32(pts + (size_t)c * n * 3, n, s, start, out + (size_t)c * s, scratch);
    free(scratch);
}

/* PCA/pointnet/PointNet.py:691-706 — centroid (double accumulate, see oracle/pointnet_numpy.py),
 * centre, max distance, scale = max(maxdist, 1e-7), divide. */
void oracle_normalize_f32(const float *x, int b, int64_t n, float *y, float *centroid, float *scale)
{
    for (int c = 0; c < b; ++c) {
        const float *p = x + (size_t)c * n * 3;
        float *q = y + (size_t)c * n * 3;
        double sx = 0, sy = 0, sz = 0;
        for (int64_t i = 0; i < n; ++i) { sx += p[3 * i]; sy += p[3 * i + 1]; sz += p[3 * i + 2]; }
        const float cx = (float)(sx / (double)n), cy = (float)(sy / (double)n), cz = (float)(sz / (double)n);
        float md = 0.0f;
        for (int64_t i = 0; i < n; ++i) {
            const float dx = p[3 * i] - cx, dy = p[3 * i + 1] - cy, dz = p[3 * i + 2] - cz;
            float dd = dx * dx + dy * dy;
            dd = dd + dz * dz;
            const float dist = sqrtf(dd);
            if (dist > md) md = dist;
        }
        const float sc = md > 1e-7f ? md : 1e-7f;
        for (int64_t i = 0; i < n; ++i) {
            q[3 * i] = (p[3 * i] - cx) / sc;
            q[3 * i + 1] = (p[3 * i + 1] - cy) / sc;
            q[3 * i + 2] = (p[3 * i + 2] - cz) / sc;
        }
        if (centroid) { centroid[3 * c] = cx; centroid[3 * c + 1] = cy; centroid[3 * c + 2] = cz; }
        if (scale) scale[c] = sc;
    }
}

typedef struct { uint64_t key; int64_t idx; } keyidx_t;
static int cmp_keyidx(const void *a, const void *b)
{
    const keyidx_t *x = (const keyidx_t *)a, *y = (const keyidx_t *)b;
    if (x->key != y->key) return x->key < y->key ? -1 : 1;
    return x->idx < y->idx ? -1 : (x->idx > y->idx);
}

/* oracle/sampling.py:voxel_downsample for one cloud.  origin == NULL -> per-axis minimum.
 * Returns the voxel count V (or -1 if a voxel index is out of range); writes V centroids, counts. */
int64_t oracle_voxel_f32(const float *pts, int64_t n, const float *origin, float voxel,
                         float *out_xyz, int32_t *out_count)
{
    float o[3];
    if (origin) { o[0] = origin[0]; o[1] = origin[1]; o[2] = origin[2]; }
    else {
        o[0] = o[1] = o[2] = INFINITY;
        for (int64_t i = 0; i < n; ++i)
            for (int a = 0; a < 3; ++a) if (pts[3 * i + a] < o[a]) o[a] = pts[3 * i + a];
    }
    keyidx_t *ki = (keyidx_t *)malloc(sizeof(keyidx_t) * (size_t)n);
    for (int64_t i = 0; i < n; ++i) {
        uint64_t key = 0;
        for (int a = 0; a < 3; ++a) {
            const float q = (pts[3 * i + a] - o[a]) / voxel;
            const int64_t k = (int64_t)floorf(q);
            if (k <= -(1 << 20) || k >= (1 << 20)) { free(ki); return -1; }
            key = (key << 21) | (uint64_t)(k + (1 << 20));
        }
        ki[i].key = key; ki[i].idx = i;
    }
    qsort(ki, (size_t)n, sizeof(keyidx_t), cmp_keyidx);
    int64_t v = 0;
    for (int64_t i = 0; i < n;) {
        int64_t j = i;
        double sx = 0, sy = 0, sz = 0;
        while (j < n && ki[j].key == ki[i].key) {
            const float *p = pts + 3 * ki[j].idx;
            sx += p[0]; sy += p[1]; sz += p[2];
            ++j;
        }
        const double c = (double)(j - i);
        out_xyz[3 * v] = (float)(sx / c); out_xyz[3 * v + 1] = (float)(sy / c); out_xyz[3 * v + 2] = (float)(sz / c);
        out_count[v] = (int32_t)(j - i);
        ++v; i = j;
    }
    free(ki);
    return v;
}
