/*
 * gs_next_oracle.c -- CPU restatement of the two neighbours of the rasterizer that SURVEY.md 8(f) rows 2 and 4 name:
 * the Adam step of the reference's optimizer and simple-knn's distCUDA2.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path (the CUDA library, the torch binding, the Python package) may
 * include, link or call this file.  It is the checker used by tests/ (and tools/bench_next.py's cpu_baseline leg).
 *
 * Parity pinning.
 *   gno_adam_step  : the reference's optimizer IS torch.optim.Adam (scene/gaussian_model.py:252), so the pin is torch
 *                    itself: tests/golden/ref_adam_cpu.npz holds parameters and moments produced by torch.optim.Adam on
 *                    the CPU for the reference's six groups and learning rates (generator: tests/golden/make_golden_adam.py,
 *                    which imports the reference's arguments.OptimizationParams and utils.general_utils.get_expon_lr_func).
 *                    tests/test_adam_cpu.py checks this function against them.
 *   gno_knn_mean_d2: the reference function needs a GPU (simple_knn.cu is CUDA only) and the reference ships no vectors
 *                    for it: PARITY UNPINNED against the reference binary this round.  Its result is defined exactly
 *                    (three smallest squared distances to other points, averaged), so it is checked against an
 *                    independent exact search (scipy cKDTree, float64) in tests/test_knn_cpu.py instead; a fixture from
 *                    the reference build is round 2's first GPU call (tests/golden/make_golden_knn_gpu.py).
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* ---------------------------------------------------------------------------------------------------------------
 * torch.optim.Adam, single tensor, no weight decay / amsgrad / maximize: torch/optim/adam.py (_multi_tensor_adam, the
 * non-capturable branch -- what `self.optimizer.step()` runs for CUDA parameters, train.py:173-175).  `step` is the
 * count AFTER this update.  Host scalars are doubles (Python floats in torch); tensor arithmetic is fp32 with the FMAs
 * of torch's functors (lerp: fma(w, end - self, self); addcmul: fma(value * t1, t2, self); addcdiv: fma(value, t1 / t2,
 * self)).  Compiled with -ffp-contract=off, so fmaf() marks every contraction. */
void gno_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, size_t n, double lr, double beta1,
                   double beta2, double eps, long long step) {
    const double bias_correction1 = 1.0 - pow(beta1, (double)step);
    const double bias_correction2 = 1.0 - pow(beta2, (double)step);
    const double step_size = lr / bias_correction1;
    const double bias_correction2_sqrt = sqrt(bias_correction2);
    const float w1 = (float)(1.0 - beta1), w2 = (float)(1.0 - beta2), b2 = (float)beta2;
    const float bc2s = (float)bias_correction2_sqrt, e = (float)eps, neg = (float)(-step_size);
    for (size_t i = 0; i < n; i++) {
        const float g = grad[i];
        float m = exp_avg[i], v = exp_avg_sq[i];
        m = fmaf(w1, g - m, m);                    /* exp_avg.lerp_(grad, 1 - beta1)                              */
        v = v * b2;                                /* exp_avg_sq.mul_(beta2)                                      */
        v = fmaf(w2 * g, g, v);                    /*           .addcmul_(grad, grad, value = 1 - beta2)          */
        const float denom = sqrtf(v) / bc2s + e;   /* (exp_avg_sq.sqrt() / bias_correction2_sqrt).add_(eps)       */
        param[i] = fmaf(neg, m / denom, param[i]); /* param.addcdiv_(exp_avg, denom, value = -step_size)          */
        exp_avg[i] = m;
        exp_avg_sq[i] = v;
    }
}

/* ---------------------------------------------------------------------------------------------------------------
 * SimpleKNN::knn (submodules/simple-knn/simple_knn.cu:185-221), restated with its own organisation: Morton codes over
 * the bounds (which include the origin: the reductions start from {0,0,0}, :190-199), a stable sort of (code, index),
 * boxes of BOX_SIZE consecutive points (:13, boxMinMax :73-119), and per point the two-stage search of boxMeanDist
 * (:152-183): a +-3 window in Morton order gives a rejection radius, then every box closer than it is walked. */
#define BOX_SIZE 1024 /* simple_knn.cu:13 */

static uint32_t prep_morton(uint32_t x) { /* simple_knn.cu:45-52 */
    x = (x | (x << 16)) & 0x030000FF;
    x = (x | (x << 8)) & 0x0300F00F;
    x = (x | (x << 4)) & 0x030C30C3;
    x = (x | (x << 2)) & 0x09249249;
    return x;
}

static uint32_t coord2morton(const float* c, const float* mn, const float* mx) { /* simple_knn.cu:54-61 */
    uint32_t q[3];
    for (int a = 0; a < 3; a++) {
        const float ext = mx[a] - mn[a];
        float t = ext > 0.f ? ((c[a] - mn[a]) / ext) * (float)((1 << 10) - 1) : 0.f;
        if (!(t >= 0.f)) t = 0.f;
        if (t > 1023.f) t = 1023.f;
        q[a] = prep_morton((uint32_t)t);
    }
    return q[0] | (q[1] << 1) | (q[2] << 2);
}

typedef struct { float mn[3], mx[3]; } minmax_t;

static float dist_box_point(const minmax_t* box, const float* p) { /* simple_knn.cu:121-131 */
    float diff[3] = {0.f, 0.f, 0.f};
    for (int a = 0; a < 3; a++)
        if (p[a] < box->mn[a] || p[a] > box->mx[a]) diff[a] = fminf(fabsf(p[a] - box->mn[a]), fabsf(p[a] - box->mx[a]));
    return fmaf(diff[2], diff[2], fmaf(diff[1], diff[1], diff[0] * diff[0]));
}

static void update_k_best(const float* ref, const float* point, float* knn) { /* updateKBest<3>, simple_knn.cu:136-150 */
    const float d[3] = {point[0] - ref[0], point[1] - ref[1], point[2] - ref[2]};
    /* FMUL, FFMA, FFMA in the reference's sm_100 SASS */
    float dist = fmaf(d[2], d[2], fmaf(d[1], d[1], d[0] * d[0]));
    for (int j = 0; j < 3; j++) {
        if (knn[j] > dist) {
            const float t = knn[j];
            knn[j] = dist;
            dist = t;
        }
    }
}

typedef struct { uint32_t code, idx; } code_idx_t;

/* cub::DeviceRadixSort::SortPairs is a STABLE sort on the key; LSD radix sort by bytes, restated */
static void stable_sort_pairs(code_idx_t* a, code_idx_t* tmp, size_t n) {
    for (int shift = 0; shift < 32; shift += 8) {
        size_t count[257];
        memset(count, 0, sizeof(count));
        for (size_t i = 0; i < n; i++) count[((a[i].code >> shift) & 255u) + 1]++;
        for (int b = 0; b < 256; b++) count[b + 1] += count[b];
        for (size_t i = 0; i < n; i++) tmp[count[(a[i].code >> shift) & 255u]++] = a[i];
        memcpy(a, tmp, n * sizeof(code_idx_t));
    }
}

int gno_knn_mean_d2(int P, const float* points /* (P,3) */, float* mean_d2 /* (P) */) {
    if (P <= 0) return 0;
    float mn[3] = {0.f, 0.f, 0.f}, mx[3] = {0.f, 0.f, 0.f}; /* init = {0,0,0}, simple_knn.cu:192 */
    for (int i = 0; i < P; i++)
        for (int a = 0; a < 3; a++) {
            mn[a] = fminf(mn[a], points[3 * (size_t)i + a]);
            mx[a] = fmaxf(mx[a], points[3 * (size_t)i + a]);
        }
    code_idx_t* order = (code_idx_t*)malloc(2 * (size_t)P * sizeof(code_idx_t));
    if (!order) return -1;
    for (int i = 0; i < P; i++) {
        order[i].code = coord2morton(points + 3 * (size_t)i, mn, mx);
        order[i].idx = (uint32_t)i; /* thrust::sequence, :206 */
    }
    stable_sort_pairs(order, order + P, (size_t)P);
    const int num_boxes = (P + BOX_SIZE - 1) / BOX_SIZE;
    minmax_t* boxes = (minmax_t*)malloc((size_t)num_boxes * sizeof(minmax_t));
    if (!boxes) { free(order); return -1; }
    for (int b = 0; b < num_boxes; b++) { /* boxMinMax, :73-119 */
        minmax_t me = {{FLT_MAX, FLT_MAX, FLT_MAX}, {-FLT_MAX, -FLT_MAX, -FLT_MAX}};
        const int end = (b + 1) * BOX_SIZE < P ? (b + 1) * BOX_SIZE : P;
        for (int i = b * BOX_SIZE; i < end; i++)
            for (int a = 0; a < 3; a++) {
                const float v = points[3 * (size_t)order[i].idx + a];
                me.mn[a] = fminf(me.mn[a], v);
                me.mx[a] = fmaxf(me.mx[a], v);
            }
        boxes[b] = me;
    }
#pragma omp parallel for schedule(dynamic, 256)
    for (int idx = 0; idx < P; idx++) { /* boxMeanDist, :152-183 */
        const float* point = points + 3 * (size_t)order[idx].idx;
        float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
        const int lo = idx - 3 > 0 ? idx - 3 : 0, hi = idx + 3 < P - 1 ? idx + 3 : P - 1;
        for (int i = lo; i <= hi; i++) {
            if (i == idx) continue;
            update_k_best(point, points + 3 * (size_t)order[i].idx, best);
        }
        const float reject = best[2];
        best[0] = best[1] = best[2] = FLT_MAX;
        for (int b = 0; b < num_boxes; b++) {
            const float dist = dist_box_point(&boxes[b], point);
            if (dist > reject || dist > best[2]) continue;
            const int end = (b + 1) * BOX_SIZE < P ? (b + 1) * BOX_SIZE : P;
            for (int i = b * BOX_SIZE; i < end; i++) {
                if (i == idx) continue;
                update_k_best(point, points + 3 * (size_t)order[i].idx, best);
            }
        }
        mean_d2[order[idx].idx] = (best[0] + best[1] + best[2]) / 3.0f; /* :182 */
    }
    free(boxes);
    free(order);
    return 0;
}
