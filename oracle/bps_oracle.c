/*
 * TEST INFRASTRUCTURE — CPU oracle for the Basis-Point-Set stage.  Never on the product path.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may call it.
 *
 * Restates (the arithmetic lives in third-party code that is NOT under /root/reference and is installed
 * unpinned from git HEAD, README.md:39,41):
 *   - bps_torch.bps.bps_torch.encode  (call sites DexGanGrasp/data/bps_encoder.py:19,33;
 *     DexGanGrasp/utils/utils.py:134-136; dexgangrasp_real_robot.py:101,139): x -> float32, per cloud
 *     nearest cloud point to every basis point, deltas = x[idx] - basis, dists = sqrt(sum(deltas^2)).
 *   - chamfer_distance NmDistanceKernel (otaheri/chamfer_distance): fp32 direct differences,
 *     d = dx*dx + dy*dy + dz*dz as nvcc contracts it (fma(dz,dz, fma(dy,dy, dx*dx))), strict `<`
 *     so the lowest point index wins a tie.
 * Pinned by the reference's own fixtures: data/real_objects/object/<name>.pcd + models/basis_point_set.npy
 * must give data/real_objects/bps/<name>.npy bit-exactly; that holds only when the final sum is evaluated
 * as (dx^2 + dz^2) + dy^2 with separately rounded squares (tests/test_oracle_bps.py).
 *
 * Build: see oracle/Makefile (gcc -O2 -ffp-contract=off -pthread).  No libgomp in the image, so the
 * multi-core entry point splits the basis range over pthreads.
 */
#include <math.h>
#include <stdint.h>
#include <stddef.h>
#include <pthread.h>

/* squared distance used for the arg-min search (chamfer kernel form) */
static inline float search_d2(float bx, float by, float bz, float x, float y, float z) {
    float dx = x - bx, dy = y - by, dz = z - bz;
    return fmaf(dz, dz, fmaf(dy, dy, dx * dx));
}

/* distance written to the encoding (bps_torch post-ops as torch's CUDA sum(dim=2) evaluated them) */
static inline float encode_dist(float bx, float by, float bz, float x, float y, float z) {
    float dx = x - bx, dy = y - by, dz = z - bz;
    float sx = dx * dx, sy = dy * dy, sz = dz * dz; /* -ffp-contract=off: three rounded squares */
    float s = (sx + sz) + sy;
    return sqrtf(s);
}

/*
 * One cloud, basis range [p0,p1).  basis: [P,3] f32 AoS, pts: [N,3] f32 AoS.
 * dists: [P] f32, idx: [P] i32.  N == 0 -> dists = +inf, idx = -1 (same convention as the CUDA path).
 */
static void encode_range(const float* basis, int p0, int p1, const float* pts, int N, float* dists, int32_t* idx) {
    for (int p = p0; p < p1; ++p) {
        const float bx = basis[3 * p], by = basis[3 * p + 1], bz = basis[3 * p + 2];
        float best = INFINITY;
        int32_t bi = -1;
        for (int k = 0; k < N; ++k) {
            float d = search_d2(bx, by, bz, pts[3 * k], pts[3 * k + 1], pts[3 * k + 2]);
            if (bi < 0 || d < best) { best = d; bi = k; }
        }
        idx[p] = bi;
        dists[p] = bi < 0 ? INFINITY : encode_dist(bx, by, bz, pts[3 * bi], pts[3 * bi + 1], pts[3 * bi + 2]);
    }
}

void dexgg_oracle_bps_encode(const float* basis, int P, const float* pts, int N, float* dists, int32_t* idx) {
    encode_range(basis, 0, P, pts, N, dists, idx);
}

typedef struct {
    const float* basis; int p0, p1; const float* pts; int N; float* dists; int32_t* idx;
} job_t;

static void* job_main(void* a) {
    job_t* j = (job_t*)a;
    encode_range(j->basis, j->p0, j->p1, j->pts, j->N, j->dists, j->idx);
    return NULL;
}

/* Same result as dexgg_oracle_bps_encode, basis range split over `nthreads` pthreads (<= 256). */
void dexgg_oracle_bps_encode_mt(const float* basis, int P, const float* pts, int N, float* dists, int32_t* idx,
                                int nthreads) {
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 256) nthreads = 256;
    pthread_t th[256];
    job_t jobs[256];
    int per = (P + nthreads - 1) / nthreads;
    int used = 0;
    for (int t = 0; t < nthreads; ++t) {
        int p0 = t * per, p1 = p0 + per > P ? P : p0 + per;
        if (p0 >= p1) break;
        jobs[t] = (job_t){basis, p0, p1, pts, N, dists, idx};
        pthread_create(&th[t], NULL, job_main, &jobs[t]);
        ++used;
    }
    for (int t = 0; t < used; ++t) pthread_join(th[t], NULL);
}

/* Ragged batch: offsets[B+1] are point offsets into pts; outputs are [B,P]. */
void dexgg_oracle_bps_encode_batch(const float* basis, int P, const float* pts, const int64_t* offsets, int B,
                                   float* dists, int32_t* idx, int nthreads) {
    for (int b = 0; b < B; ++b) {
        int64_t o = offsets[b];
        int N = (int)(offsets[b + 1] - o);
        dexgg_oracle_bps_encode_mt(basis, P, pts + 3 * o, N, dists + (size_t)b * P, idx + (size_t)b * P, nthreads);
    }
}

/* Alternative summation orders, exported so the test can show that only one reproduces the fixtures. */
void dexgg_oracle_bps_dist_variants(const float* basis, int P, const float* pts, const int32_t* idx,
                                    float* d_xz_y, float* d_xy_z, float* d_fma) {
    for (int p = 0; p < P; ++p) {
        const float bx = basis[3 * p], by = basis[3 * p + 1], bz = basis[3 * p + 2];
        int k = idx[p];
        float dx = pts[3 * k] - bx, dy = pts[3 * k + 1] - by, dz = pts[3 * k + 2] - bz;
        float sx = dx * dx, sy = dy * dy, sz = dz * dz;
        d_xz_y[p] = sqrtf((sx + sz) + sy);
        d_xy_z[p] = sqrtf((sx + sy) + sz);
        d_fma[p] = sqrtf(fmaf(dz, dz, fmaf(dy, dy, dx * dx)));
    }
}
