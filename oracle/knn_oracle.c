/*
 * TEST INFRASTRUCTURE ONLY -- CPU restatement of simple-knn's distCUDA2.
 * Follows KNN/simple_knn.cu:119-183 (KNN = submodules/simple-knn of KTH-RPL/cloth-splatting):
 * for every point the mean of the three smallest squared distances to the OTHER indices
 * (duplicates count, distance 0).  The reference's Morton/box pruning is conservative
 * (simple_knn.cu:167-170), so its result equals this exhaustive search; fewer than 4 points leave
 * FLT_MAX entries in the sum exactly like the reference's initial best[] (simple_knn.cu:154).
 * Parity status: pinned against oracle/_ref/libref_cuda.so outputs in tests/golden/.
 */
#include <float.h>
#include <math.h>
#include <stdint.h>

void oracle_dist2(int P, const float* pts, float* out) {
#pragma omp parallel for schedule(dynamic, 64)
    for (int i = 0; i < P; ++i) {
        const float rx = pts[3 * i], ry = pts[3 * i + 1], rz = pts[3 * i + 2];
        float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
        for (int j = 0; j < P; ++j) {
            if (j == i) continue;
            const float dx = pts[3 * j] - rx, dy = pts[3 * j + 1] - ry, dz = pts[3 * j + 2] - rz;
            float dist = fmaf(dz, dz, fmaf(dx, dx, dy * dy)); /* nvcc's contraction of dx*dx + dy*dy + dz*dz (reference SASS) */
            for (int k = 0; k < 3; ++k)
                if (best[k] > dist) { const float t = best[k]; best[k] = dist; dist = t; }
        }
        out[i] = (best[0] + best[1] + best[2]) / 3.0f;
    }
}
