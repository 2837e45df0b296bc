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
 *   gno_knn_mean_d2: the reference function needs a GPU (simple_knn.cu is CUDA only) and ships no vectors.  PINNED since
 *                    round 2: tests/golden/ref_knn_cuda.npz was written on a B200 by the reference's own simple-knn built
 *                    for sm_100 (oracle/build_ref.py: build_knn; generator tests/golden/make_golden_knn_gpu.py), and this
 *                    function reproduces all six cases bit for bit (tests/test_knn_cpu.py).  The fixture is what showed
 *                    that nvcc contracts d.x*d.x + d.y*d.y + d.z*d.z as fma(dz, dz, fma(dx, dx, dy*dy)).  Also checked
 *                    against an independent exact search (scipy cKDTree, float64).
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
    return fmaf(diff[2], diff[2], fmaf(diff[0], diff[0], diff[1] * diff[1])); /* FMUL y, FFMA x, FFMA z in the reference's sm_100 SASS */
}

static void update_k_best(const float* ref, const float* point, float* knn) { /* updateKBest<3>, simple_knn.cu:136-150 */
    const float d[3] = {point[0] - ref[0], point[1] - ref[1], point[2] - ref[2]};
    /* the reference's sm_100 SASS: FMUL on the y difference, then FFMA x, then FFMA z */
    float dist = fmaf(d[2], d[2], fmaf(d[0], d[0], d[1] * d[1]));
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

/* ===============================================================================================================
 * Densification (second half): the reference's OWN sequence, round by round, on growing arrays --
 * deliberately not the product's one-pass formulation, so that agreement between the two checks the collapse of the
 * three rounds into per-original decisions.
 * Parity pinning: tests/golden/ref_densify_cpu.npz holds inputs, the recorded torch.normal samples and the outputs of the
 * reference's unmodified GaussianModel.densify_and_prune / add_densification_stats (scene/gaussian_model.py:558-602), run
 * on the CPU by tests/golden/make_golden_densify.py (the reference hard-codes device="cuda"; the script maps that device
 * name to the CPU and changes nothing else).  tests/test_densify_cpu.py checks these functions against that file:
 * selections, order and copied values exactly, computed values (child xyz / scaling) to 1e-6.
 * Arithmetic notes: torch's CUDA division by a Python scalar multiplies by the float32 reciprocal (BinaryDivTrueKernel.cu),
 * restated here as exp(s) * (1.0f / (float)(0.8 N)); the CPU run of the reference divides, which differs by at most 1 ulp.
 * =============================================================================================================== */
typedef struct {
    int P, ng, has_state;
    int w[8];
    float* p[8]; float* m[8]; float* v[8];
} gno_model_t;

static float gno_sigmoid(float x) { return 1.0f / (1.0f + expf(-x)); }
static float gno_max3(const float* s) { float a = expf(s[0]), b = expf(s[1]), c = expf(s[2]); float m = a > b ? a : b; return m > c ? m : c; }

/* rows with mask != 0 survive: _prune_optimizer (scene/gaussian_model.py:394-419) */
static void gno_select(gno_model_t* M, const unsigned char* keep) {
    int n = 0;
    for (int i = 0; i < M->P; i++) n += keep[i] != 0;
    for (int g = 0; g < M->ng; g++) {
        const int w = M->w[g];
        float** arrs[3] = {&M->p[g], &M->m[g], &M->v[g]};
        for (int a = 0; a < (M->has_state ? 3 : 1); a++) {
            float* src = *arrs[a];
            float* dst = (float*)malloc(sizeof(float) * (size_t)(n > 0 ? n : 1) * w);
            size_t o = 0;
            for (int i = 0; i < M->P; i++)
                if (keep[i]) { memcpy(dst + o, src + (size_t)i * w, sizeof(float) * w); o += w; }
            free(src);
            *arrs[a] = dst;
        }
    }
    M->P = n;
}

/* cat_tensors_to_optimizer (scene/gaussian_model.py:445-470): parameters extended, moments extended with zeros */
static void gno_cat(gno_model_t* M, float* const* ext, int n_new) {
    for (int g = 0; g < M->ng; g++) {
        const size_t w = M->w[g], old = (size_t)M->P * w, add = (size_t)n_new * w;
        M->p[g] = (float*)realloc(M->p[g], sizeof(float) * (old + add + 1));
        memcpy(M->p[g] + old, ext[g], sizeof(float) * add);
        if (M->has_state) {
            M->m[g] = (float*)realloc(M->m[g], sizeof(float) * (old + add + 1));
            M->v[g] = (float*)realloc(M->v[g], sizeof(float) * (old + add + 1));
            memset(M->m[g] + old, 0, sizeof(float) * add);
            memset(M->v[g] + old, 0, sizeof(float) * add);
        }
    }
    M->P += n_new;
}

/* build_rotation (utils/general_utils.py:90-114) then bmm with one sample and the parent's position (gaussian_model.py:519-520) */
static void gno_child_xyz(const float* r4, const float* xyz, const float* smp, float* out) {
    const float n = sqrtf(r4[0] * r4[0] + r4[1] * r4[1] + r4[2] * r4[2] + r4[3] * r4[3]);
    const float r = r4[0] / n, x = r4[1] / n, y = r4[2] / n, z = r4[3] / n;
    const float R[9] = {1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                        2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                        2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)};
    for (int c = 0; c < 3; c++)
        out[c] = fmaf(R[3 * c + 2], smp[2], fmaf(R[3 * c + 1], smp[1], R[3 * c] * smp[0])) + xyz[c];
}

/* The two selection masks of densify_and_clone / densify_and_split for the P original Gaussians (the split mask is
 * evaluated by the reference on the P + n_clone rows after cloning, with zero gradients for the clones: none of those can
 * be selected).  Returns n_split; the caller draws torch.normal(0, exp(scaling)[split].repeat(N,1)) from it. */
int gno_densify_masks(int P, const float* accum, const float* denom, const float* scaling, double max_grad, double extent,
                      double percent_dense, unsigned char* clone_mask, unsigned char* split_mask) {
    const float thr = (float)max_grad, dense = (float)(percent_dense * extent);
    int ns = 0;
    for (int i = 0; i < P; i++) {
        float g = accum[i] / denom[i];
        if (isnan(g)) g = 0.f;
        const float smax = gno_max3(scaling + 3 * (size_t)i);
        clone_mask[i] = sqrtf(g * g) >= thr && smax <= dense;
        split_mask[i] = g >= thr && smax > dense;
        ns += split_mask[i];
    }
    return ns;
}

/* densify_and_prune (scene/gaussian_model.py:558-597).  params / exp_avg / exp_avg_sq: ng arrays of (P, widths[g]);
 * exp_avg may be NULL (optimizer without state).  Group roles by index.  samples: (N * n_split, 3) as gno_densify_masks
 * counted them.  Outputs are written into caller arrays of `capacity` rows (P + n_clone + N * n_split always suffices).
 * Returns the new number of Gaussians, or -1 when capacity is too small. */
int gno_densify_and_prune(int P, int ng, const int* widths, const float* const* params, const float* const* exp_avg,
                          const float* const* exp_avg_sq, const float* accum, const float* denom, double max_grad,
                          double min_opacity, double extent, double percent_dense, double max_screen_size, int N,
                          const float* samples, int xyz_g, int scaling_g, int rotation_g, int opacity_g, int capacity,
                          float* const* params_out, float* const* exp_avg_out, float* const* exp_avg_sq_out) {
    gno_model_t M;
    M.P = P; M.ng = ng; M.has_state = exp_avg != NULL;
    for (int g = 0; g < ng; g++) {
        const size_t n = (size_t)P * widths[g];
        M.w[g] = widths[g];
        M.p[g] = (float*)malloc(sizeof(float) * (n + 1)); memcpy(M.p[g], params[g], sizeof(float) * n);
        M.m[g] = M.v[g] = NULL;
        if (M.has_state) {
            M.m[g] = (float*)malloc(sizeof(float) * (n + 1)); memcpy(M.m[g], exp_avg[g], sizeof(float) * n);
            M.v[g] = (float*)malloc(sizeof(float) * (n + 1)); memcpy(M.v[g], exp_avg_sq[g], sizeof(float) * n);
        }
    }
    const float thr = (float)max_grad, dense = (float)(percent_dense * extent);
    float* grads = (float*)malloc(sizeof(float) * (size_t)(P + 1));
    for (int i = 0; i < P; i++) { float g = accum[i] / denom[i]; grads[i] = isnan(g) ? 0.f : g; }          /* :566-567 */

    /* ---- densify_and_clone (:538-556) ---- */
    {
        unsigned char* sel = (unsigned char*)calloc((size_t)P + 1, 1);
        int n = 0;
        for (int i = 0; i < P; i++) { sel[i] = sqrtf(grads[i] * grads[i]) >= thr && gno_max3(M.p[scaling_g] + 3 * (size_t)i) <= dense; n += sel[i]; }
        float* ext[8];
        for (int g = 0; g < ng; g++) {
            const int w = M.w[g];
            ext[g] = (float*)malloc(sizeof(float) * ((size_t)n * w + 1));
            size_t o = 0;
            for (int i = 0; i < P; i++) if (sel[i]) { memcpy(ext[g] + o, M.p[g] + (size_t)i * w, sizeof(float) * w); o += w; }
        }
        gno_cat(&M, ext, n);
        for (int g = 0; g < ng; g++) free(ext[g]);
        free(sel);
    }

    /* ---- densify_and_split (:497-536) ---- */
    {
        const int n_init = M.P;
        float* padded = (float*)calloc((size_t)n_init + 1, sizeof(float));
        memcpy(padded, grads, sizeof(float) * (size_t)P);
        unsigned char* sel = (unsigned char*)calloc((size_t)n_init + 1, 1);
        int ns = 0;
        for (int i = 0; i < n_init; i++) { sel[i] = padded[i] >= thr && gno_max3(M.p[scaling_g] + 3 * (size_t)i) > dense; ns += sel[i]; }
        int* idx = (int*)malloc(sizeof(int) * (size_t)(ns + 1));
        for (int i = 0, j = 0; i < n_init; i++) if (sel[i]) idx[j++] = i;
        const int n_new = N * ns;
        const float inv = 1.0f / (float)(0.8 * N);
        float* ext[8];
        for (int g = 0; g < ng; g++) {
            const int w = M.w[g];
            ext[g] = (float*)malloc(sizeof(float) * ((size_t)n_new * w + 1));
            for (int j = 0; j < n_new; j++) {                                  /* .repeat(N, 1): row j comes from parent j mod ns */
                const int par = idx[j % (ns > 0 ? ns : 1)];
                float* d = ext[g] + (size_t)j * w;
                if (g == xyz_g) gno_child_xyz(M.p[rotation_g] + 4 * (size_t)par, M.p[xyz_g] + 3 * (size_t)par, samples + 3 * (size_t)j, d);
                else if (g == scaling_g) for (int c = 0; c < 3; c++) d[c] = logf(expf(M.p[g][3 * (size_t)par + c]) * inv);
                else memcpy(d, M.p[g] + (size_t)par * w, sizeof(float) * w);
            }
        }
        gno_cat(&M, ext, n_new);
        for (int g = 0; g < ng; g++) free(ext[g]);
        unsigned char* keep = (unsigned char*)malloc((size_t)M.P + 1);         /* prune_filter = cat(selected, zeros) -> keep = ~filter */
        for (int i = 0; i < M.P; i++) keep[i] = !(i < n_init && sel[i]);
        gno_select(&M, keep);
        free(keep); free(idx); free(sel); free(padded);
    }

    /* ---- final prune (:577-593); max_radii2D was reset to zeros by densification_postfix (:493-495) ---- */
    {
        const float min_op = (float)min_opacity, big = (float)(0.1 * extent), mss = (float)max_screen_size;
        unsigned char* keep = (unsigned char*)malloc((size_t)M.P + 1);
        for (int i = 0; i < M.P; i++) {
            int prune = gno_sigmoid(M.p[opacity_g][i]) < min_op;
            if (max_screen_size != 0.0) {
                const float max_radii2D = 0.f;
                prune = prune || max_radii2D > mss || gno_max3(M.p[scaling_g] + 3 * (size_t)i) > big;
            }
            keep[i] = !prune;
        }
        gno_select(&M, keep);
        free(keep);
    }
    free(grads);
    int ret = M.P;
    if (M.P > capacity) ret = -1;
    for (int g = 0; g < ng; g++) {
        const size_t n = (size_t)M.P * M.w[g];
        if (ret >= 0) {
            memcpy(params_out[g], M.p[g], sizeof(float) * n);
            if (M.has_state) { memcpy(exp_avg_out[g], M.m[g], sizeof(float) * n); memcpy(exp_avg_sq_out[g], M.v[g], sizeof(float) * n); }
        }
        free(M.p[g]); free(M.m[g]); free(M.v[g]);
    }
    return ret;
}

/* train.py:153-158 + add_densification_stats (scene/gaussian_model.py:599-602), visibility = radii > 0 */
void gno_densify_stats(int P, const float* vs_grad /* (P,3) */, const int* radii, float* accum, float* denom, float* max_radii2D) {
    for (int i = 0; i < P; i++) {
        if (radii[i] <= 0) continue;
        const float gx = vs_grad[3 * (size_t)i], gy = vs_grad[3 * (size_t)i + 1];
        if (max_radii2D) max_radii2D[i] = fmaxf(max_radii2D[i], (float)radii[i]);
        accum[i] += sqrtf(fmaf(gy, gy, gx * gx));
        denom[i] += 1.0f;
    }
}
