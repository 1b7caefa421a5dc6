/*
 * rm_oracle.c — CPU restatement of the reference's per-pixel sphere-tracing pass.
 *
 * TEST INFRASTRUCTURE ONLY.  This file is the *checker* for the CUDA path and the timed
 * CPU baseline of bench.py.  Nothing under ray-marching_b200/ links, loads or calls it; only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may.
 *
 * PARITY UNPINNED: the reference (Mesoptier/ray-marching) ships no tests, golden images or
 * known-answer vectors for this path, and it cannot be built or run here
 * (Rust + WGSL through wgpu; no rustc/cargo/naga, no graphics stack — SURVEY.md F8/F9).  The
 * oracle is therefore pinned only by the known-answer tests K1–K8 that tests/ derives from
 * the shader text, and by an independent numpy restatement (oracle/np_oracle.py).
 *
 * What it follows, line by line (all citations relative to /root/reference):
 *   src/ray_marching/ray_marching.wgsl:36-76    fs_main  (16-sample AA, ray gen, gamma, mean)
 *   src/ray_marching/ray_marching.wgsl:87-131   ray_march (march loop, hit shade, floor, sky)
 *   src/ray_marching/ray_marching.wgsl:135-144  calculate_normal (tetrahedron taps)
 *   src/ray_marching/ray_marching.wgsl:146-252  command buffer cursor, value stack, map_scene,
 *                                               eval_cmd_{sphere,box,union,subtract}
 *   src/ray_marching/renderer.rs:29-41          Uniforms (144 B) / RayMarchLimits (12 B) layouts
 *   src/ray_marching/csg/builder.rs:1-62        wire format of the u32 word buffer
 * plus the two build-defined extension opcodes of SURVEY.md §8(a) X1/X2 (Intersection = 102,
 * SmoothUnion = 103 [k]) which have no counterpart in the reference, and the N4 extension
 *  at the opcode values the reference's enum reserves in comments
 * (builder.rs:8,16-23): Plane = 2 and the space-transform pairs TranslationPush/Pop = 200/201,
 * RotationPush/Pop = 202/203, ScalePush/Pop = 204/205, evaluated on a position stack the way
 * the reserved names describe ("1 child, transforms space").  Normative semantics: map_scene().
 *
 * Numerics: float32 throughout, IEEE add/mul/div/sqrt, NO fused multiply-add (build with
 * -ffp-contract=off), evaluation order exactly as SURVEY.md §8(c).1 rules 1–11.
 *   min(a,b) = (b < a) ? b : a      max(a,b) = (a < b) ? b : a     (WGSL wording; for NaN / ±0
 *   operands the result may differ in sign-of-zero from other conformant implementations —
 *   irrelevant for every comparison made downstream.)
 *   round = rintf under the default RNE mode (Q7); f32→i32 saturating, NaN→0 (Q6).
 *   unorm8 = floor(clamp(v,0,1)*255 + 0.5).
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define RMO_OK 0
#define RMO_E_INVALID_ARG (-1)
#define RMO_E_BAD_PROGRAM (-2)

#define RMO_STACK_MAX 32 /* ray_marching.wgsl:173 */
#define RMO_POS_STACK_MAX 8 /* N4: nesting depth of space transforms (build-defined) */

/* sample classes written to the per-sample AOV */
#define RMO_CLS_SKY 0
#define RMO_CLS_FLOOR 1
#define RMO_CLS_OBJECT 2

typedef struct {
    float extent[2]; /* viewport_extent @0          wgsl:24 */
    float pad_[2];   /* mat4x4 is 16-byte aligned           */
    float inv_proj[16]; /* column-major @16          wgsl:26 */
    float inv_view[16]; /* column-major @80          wgsl:28 */
} rmo_uniforms; /* 144 bytes — renderer.rs:29-34 via encase UniformBuffer */

typedef struct {
    float min_dist;    /* wgsl:79 */
    float max_dist;    /* wgsl:80 */
    uint32_t max_iter; /* wgsl:81 */
} rmo_limits; /* 12 bytes — renderer.rs:36-41 */

typedef struct {
    uint32_t cmd_count;
    const uint32_t *words;
    uint32_t n_words;
    rmo_limits lim;
} rmo_scene;

static inline float f_min(float a, float b) { return (b < a) ? b : a; }
static inline float f_max(float a, float b) { return (a < b) ? b : a; }
static inline float f_bits(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

/* matrix (column-major) * vec4, as the sum of columns scaled by the vector components,
 * accumulated left to right (SURVEY.md §8(c).1 rule 3). */
static inline void mat4_mul_vec4(const float *m, const float v[4], float out[4]) {
    for (int r = 0; r < 4; ++r) {
        float acc = m[0 + r] * v[0];
        acc = acc + m[4 + r] * v[1];
        acc = acc + m[8 + r] * v[2];
        acc = acc + m[12 + r] * v[3];
        out[r] = acc;
    }
}

/* ray_marching.wgsl:187-252 (+ X1/X2).  Caller guarantees the program was validated. */
static float map_scene(const rmo_scene *sc, float px, float py, float pz) {
    if (sc->cmd_count == 0u) return sc->lim.max_dist; /* wgsl:189-191 */
    float stack[RMO_STACK_MAX];
    uint32_t sp = 0;   /* wgsl:194 */
    uint32_t ptr = 0;  /* wgsl:195 */
    const uint32_t *w = sc->words;
    float pos_stack[RMO_POS_STACK_MAX][3]; /* N4: saved evaluation points */
    uint32_t psp = 0;
    for (uint32_t idx = 0; idx < sc->cmd_count; ++idx) { /* wgsl:197 */
        uint32_t type = w[ptr++];                          /* wgsl:198 */
        float v;
        switch (type) {
        /* ---- N4 space transforms: no value is pushed; Push saves the evaluation point and
         * replaces it, Pop restores it (builder.rs:16-23 "1 child, transforms space") ---- */
        case 200u: { /* TranslationPush [200, tx,ty,tz]: pos = pos - t */
            float tx = f_bits(w[ptr]), ty = f_bits(w[ptr + 1]), tz = f_bits(w[ptr + 2]);
            ptr += 3;
            pos_stack[psp][0] = px; pos_stack[psp][1] = py; pos_stack[psp][2] = pz; ++psp;
            px = px - tx; py = py - ty; pz = pz - tz;
            continue;
        }
        case 202u: { /* RotationPush [202, m00,m01,m02, m10,m11,m12, m20,m21,m22] (row-major R,
                      * object -> world): pos = transpose(R) * pos, each row summed left to right */
            const float *m = (const float *)(const void *)(w + ptr);
            float M[9];
            memcpy(M, m, sizeof M);
            ptr += 9;
            pos_stack[psp][0] = px; pos_stack[psp][1] = py; pos_stack[psp][2] = pz; ++psp;
            float x = (M[0] * px + M[3] * py) + M[6] * pz;
            float y = (M[1] * px + M[4] * py) + M[7] * pz;
            float z = (M[2] * px + M[5] * py) + M[8] * pz;
            px = x; py = y; pz = z;
            continue;
        }
        case 204u: { /* ScalePush [204, s]: pos = pos / s (uniform scale, s > 0) */
            float s = f_bits(w[ptr]);
            ptr += 1;
            pos_stack[psp][0] = px; pos_stack[psp][1] = py; pos_stack[psp][2] = pz; ++psp;
            px = px / s; py = py / s; pz = pz / s;
            continue;
        }
        case 201u: case 203u: /* TranslationPop [201], RotationPop [203]: restore pos */
            --psp;
            px = pos_stack[psp][0]; py = pos_stack[psp][1]; pz = pos_stack[psp][2];
            continue;
        case 205u: { /* ScalePop [205, s]: restore pos; the child's distance is in scaled units */
            float s = f_bits(w[ptr]);
            ptr += 1;
            --psp;
            px = pos_stack[psp][0]; py = pos_stack[psp][1]; pz = pos_stack[psp][2];
            stack[sp - 1] = stack[sp - 1] * s;
            continue;
        }
        case 2u: { /* N4 plane [2, nx,ny,nz, h]: dot(pos, n) + h  (n expected unit length) */
            float nx = f_bits(w[ptr]), ny = f_bits(w[ptr + 1]), nz = f_bits(w[ptr + 2]);
            float h = f_bits(w[ptr + 3]);
            ptr += 4;
            v = ((px * nx + py * ny) + pz * nz) + h;
            break;
        }
        case 0u: { /* sphere — wgsl:229-233 */
            float cx = f_bits(w[ptr]), cy = f_bits(w[ptr + 1]), cz = f_bits(w[ptr + 2]);
            float r = f_bits(w[ptr + 3]);
            ptr += 4;
            float dx = px - cx, dy = py - cy, dz = pz - cz;
            v = sqrtf((dx * dx + dy * dy) + dz * dz) - r;
            break;
        }
        case 1u: { /* box — wgsl:235-240 */
            float cx = f_bits(w[ptr]), cy = f_bits(w[ptr + 1]), cz = f_bits(w[ptr + 2]);
            float rx = f_bits(w[ptr + 3]), ry = f_bits(w[ptr + 4]), rz = f_bits(w[ptr + 5]);
            ptr += 6;
            float qx = fabsf(px - cx) - rx, qy = fabsf(py - cy) - ry, qz = fabsf(pz - cz) - rz;
            float mx = f_max(qx, 0.0f), my = f_max(qy, 0.0f), mz = f_max(qz, 0.0f);
            float len = sqrtf((mx * mx + my * my) + mz * mz);
            v = len + f_min(f_max(qx, f_max(qy, qz)), 0.0f);
            break;
        }
        case 100u: { /* union — wgsl:242-246 */
            float b = stack[--sp];
            float a = stack[--sp];
            v = f_min(a, b);
            break;
        }
        case 101u: { /* subtraction — wgsl:248-252 */
            float b = stack[--sp];
            float a = stack[--sp];
            v = f_max(a, -b);
            break;
        }
        case 102u: { /* X1 intersection (extension) */
            float b = stack[--sp];
            float a = stack[--sp];
            v = f_max(a, b);
            break;
        }
        case 103u: { /* X2 smooth union (extension): opcode then param k (builder.rs:39-40 order) */
            float k = f_bits(w[ptr]);
            ptr += 1;
            float b = stack[--sp];
            float a = stack[--sp];
            float h = f_max(k - fabsf(a - b), 0.0f) / k;
            v = f_min(a, b) - ((h * h) * k) * 0.25f;
            break;
        }
        default: /* wgsl:223-225; rejected by validation, never reached */
            v = 0.0f;
            break;
        }
        stack[sp++] = v; /* wgsl:199 */
    }
    return stack[--sp]; /* wgsl:202 */
}

/* Static check of the word stream: opcode set, word count, stack discipline.  Mirrors the
 * host-side validation of the product so the oracle never reads out of bounds. */
static int validate_program(uint32_t cmd_count, const uint32_t *w, uint32_t n_words) {
    uint32_t ptr = 0;
    int sp = 0;
    /* N4: open transform groups — opcode of the push, value-stack depth at the push, scale bits */
    uint32_t g_op[RMO_POS_STACK_MAX], g_s[RMO_POS_STACK_MAX];
    int g_sp[RMO_POS_STACK_MAX];
    int psp = 0;
    if (cmd_count == 0) return RMO_OK;
    if (!w) return RMO_E_INVALID_ARG;
    for (uint32_t i = 0; i < cmd_count; ++i) {
        if (ptr >= n_words) return RMO_E_BAD_PROGRAM;
        uint32_t t = w[ptr++];
        uint32_t np;
        int pops;
        switch (t) {
        case 0u: np = 4; pops = 0; break;
        case 1u: np = 6; pops = 0; break;
        case 2u: np = 4; pops = 0; break;
        case 100u: case 101u: case 102u: np = 0; pops = 2; break;
        case 103u: np = 1; pops = 2; break;
        case 200u: np = 3; pops = -1; break;
        case 202u: np = 9; pops = -1; break;
        case 204u: np = 1; pops = -1; break;
        case 201u: case 203u: np = 0; pops = -2; break;
        case 205u: np = 1; pops = -2; break;
        default: return RMO_E_BAD_PROGRAM;
        }
        if (ptr + np > n_words) return RMO_E_BAD_PROGRAM;
        /* NaN parameters are refused (the product's decoder does the same): with them min/max as `(b < a) ? b : a` and
         * as IEEE minNum/maxNum differ, and the shader's own result is NaN everywhere */
        for (uint32_t j = 0; j < np; ++j) { float f = f_bits(w[ptr + j]); if (f != f) return RMO_E_BAD_PROGRAM; }
        if (t == 103u) { float k = f_bits(w[ptr]); if (!(k > 0.0f) || k > FLT_MAX) return RMO_E_BAD_PROGRAM; }
        if (pops == -1) { /* push: opens a group that must produce exactly one value */
            if (psp >= RMO_POS_STACK_MAX) return RMO_E_BAD_PROGRAM;
            if (t == 204u) {
                float s = f_bits(w[ptr]);
                if (!(s > 0.0f) || s > FLT_MAX) return RMO_E_BAD_PROGRAM;
            }
            g_op[psp] = t; g_sp[psp] = sp; g_s[psp] = (t == 204u) ? w[ptr] : 0u;
            ++psp;
            ptr += np;
            continue;
        }
        if (pops == -2) { /* pop: closes the innermost group, which must be of the same kind */
            if (psp == 0 || g_op[psp - 1] + 1u != t || sp != g_sp[psp - 1] + 1) return RMO_E_BAD_PROGRAM;
            if (t == 205u && w[ptr] != g_s[psp - 1]) return RMO_E_BAD_PROGRAM;
            --psp;
            ptr += np;
            continue;
        }
        ptr += np;
        /* an operator inside a group may only consume values produced inside it */
        if (sp - (psp ? g_sp[psp - 1] : 0) < pops) return RMO_E_BAD_PROGRAM;
        sp -= pops;
        sp += 1;
        if (sp > RMO_STACK_MAX) return RMO_E_BAD_PROGRAM;
    }
    if (sp < 1 || psp != 0) return RMO_E_BAD_PROGRAM;
    return RMO_OK;
}

/* f32 -> i32 conversion of WGSL `vec2<i32>(round(..))`: saturating, NaN -> 0 (Q6). */
static inline int32_t sat_i32(float f) {
    if (f != f) return 0;
    if (f >= 2147483648.0f) return INT32_MAX;
    if (f <= -2147483648.0f) return INT32_MIN;
    return (int32_t)f;
}

typedef struct {
    uint64_t n_step, n_hit, n_floor, n_sample;
} rmo_counters;

/* ray_marching.wgsl:87-131.  Returns class, writes colour and (for object hits) the marched
 * distance. */
static int ray_march(const rmo_scene *sc, const float o[3], const float d[3], float col[3],
                     float *t_out, rmo_counters *cnt) {
    float dist = 0.0f; /* wgsl:88 */
    for (uint32_t it = 0; it < sc->lim.max_iter; ++it) { /* wgsl:90 */
        float px = o[0] + d[0] * dist, py = o[1] + d[1] * dist, pz = o[2] + d[2] * dist; /* :91 */
        float s = map_scene(sc, px, py, pz); /* wgsl:94 */
        cnt->n_step++;
        if (s < sc->lim.min_dist) { /* wgsl:97 */
            /* calculate_normal — wgsl:135-144; taps (+,-,-) (-,-,+) (-,+,-) (+,+,+) */
            const float eps = 0.0001f;
            static const float K[4][3] = {{1, -1, -1}, {-1, -1, 1}, {-1, 1, -1}, {1, 1, 1}};
            float n[3] = {0, 0, 0};
            for (int k = 0; k < 4; ++k) {
                float m = map_scene(sc, px + K[k][0] * eps, py + K[k][1] * eps, pz + K[k][2] * eps);
                if (k == 0) {
                    n[0] = K[k][0] * m; n[1] = K[k][1] * m; n[2] = K[k][2] * m;
                } else {
                    n[0] = n[0] + K[k][0] * m; n[1] = n[1] + K[k][1] * m; n[2] = n[2] + K[k][2] * m;
                }
            }
            float nl = sqrtf((n[0] * n[0] + n[1] * n[1]) + n[2] * n[2]);
            n[0] = n[0] / nl; n[1] = n[1] / nl; n[2] = n[2] / nl;
            /* wgsl:100-105 */
            float lx = px - 2.0f, ly = py - (-5.0f), lz = pz - 3.0f;
            float ll = sqrtf((lx * lx + ly * ly) + lz * lz);
            lx = lx / ll; ly = ly / ll; lz = lz / ll;
            float diffuse = f_max(0.02f, (n[0] * lx + n[1] * ly) + n[2] * lz);
            col[0] = 0.4f * diffuse; col[1] = 0.7f * diffuse; col[2] = 0.1f * diffuse;
            *t_out = dist;
            cnt->n_hit++;
            return RMO_CLS_OBJECT;
        }
        if (s > sc->lim.max_dist) break; /* wgsl:109-111 */
        dist = dist + s;                 /* wgsl:114 */
    }
    /* floor plane — wgsl:119-128 */
    const float floor_y = -1.5f;
    float floor_dist = (floor_y - o[1]) / d[1];
    *t_out = floor_dist;
    if (floor_dist > 0.0f) {
        float fx = o[0] + d[0] * floor_dist, fz = o[2] + d[2] * floor_dist;
        int32_t ix = sat_i32(rintf(fx + 0.5f)), iz = sat_i32(rintf(fz + 0.5f));
        float c = (float)((ix ^ iz) & 1);
        col[0] = 0.1f + 0.2f * c; col[1] = 0.1f + 0.2f * c; col[2] = 0.2f + 0.2f * c;
        cnt->n_floor++;
        return RMO_CLS_FLOOR;
    }
    col[0] = col[1] = col[2] = 0.0f; /* wgsl:130 */
    return RMO_CLS_SKY;
}

static inline uint8_t unorm8(float v) {
    float c = v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v);
    if (c != c) c = 0.0f;
    return (uint8_t)floorf(c * 255.0f + 0.5f);
}

/* fs_main for one pixel — wgsl:36-76 with the pixel→pt_screen closed form of SURVEY.md A1. */
static void shade_pixel(const rmo_uniforms *U, const rmo_scene *sc, uint32_t W, uint32_t H,
                        uint32_t px, uint32_t py, uint8_t rgba[4], uint8_t *cls16, float *t16,
                        rmo_counters *cnt) {
    float sx = ((float)px + 0.5f) / (float)W * 2.0f - 1.0f;
    float sy = 1.0f - ((float)py + 0.5f) / (float)H * 2.0f;
    const float ro_view[4] = {0.0f, 0.0f, 0.0f, 1.0f};
    float ro[4];
    mat4_mul_vec4(U->inv_view, ro_view, ro); /* wgsl:39-40 */
    float total[3] = {0, 0, 0};
    for (uint32_t i = 0; i < 4; ++i) {     /* wgsl:44 */
        for (uint32_t j = 0; j < 4; ++j) { /* wgsl:45 */
            float offx = ((float)i + 0.5f) / 4.0f - 0.5f; /* wgsl:52 */
            float offy = ((float)j + 0.5f) / 4.0f - 0.5f;
            float ax = offx / U->extent[0] * 2.0f; /* wgsl:53 */
            float ay = offy / U->extent[1] * 2.0f;
            float ndc[4] = {sx + ax, sy + ay, -1.0f, 1.0f}; /* wgsl:56-57 */
            float pv[4], pw[4];
            mat4_mul_vec4(U->inv_proj, ndc, pv); /* wgsl:58 */
            mat4_mul_vec4(U->inv_view, pv, pw);  /* wgsl:59 */
            float d4[4] = {pw[0] - ro[0], pw[1] - ro[1], pw[2] - ro[2], pw[3] - ro[3]};
            float len = sqrtf(((d4[0] * d4[0] + d4[1] * d4[1]) + d4[2] * d4[2]) + d4[3] * d4[3]);
            float rd[3] = {d4[0] / len, d4[1] / len, d4[2] / len}; /* wgsl:62 */
            float col[3], t = 0.0f;
            int c = ray_march(sc, ro, rd, col, &t, cnt); /* wgsl:65 */
            cnt->n_sample++;
            total[0] = total[0] + sqrtf(col[0]); /* wgsl:68-69 */
            total[1] = total[1] + sqrtf(col[1]);
            total[2] = total[2] + sqrtf(col[2]);
            if (cls16) cls16[i * 4 + j] = (uint8_t)c;
            if (t16) t16[i * 4 + j] = t;
        }
    }
    rgba[0] = unorm8(total[0] / 16.0f); /* wgsl:73-75 */
    rgba[1] = unorm8(total[1] / 16.0f);
    rgba[2] = unorm8(total[2] / 16.0f);
    rgba[3] = 255;
}

/*
 * Render rows [row0, row0+nrows) of a W×H frame.
 *   rgba   : nrows*W*4 bytes, row-major, first row = image row `row0` (row 0 = top).
 *   cls    : optional nrows*W*16 bytes — per-sample class (sample index i*4+j).
 *   t      : optional nrows*W*16 floats — per-sample marched distance (object hits) or
 *            floor_dist (misses).
 *   counters[4] : optional {N_step, N_hit, N_floor, N_sample}.
 *   nthreads <= 0 : all OpenMP threads.
 */
int rmo_render(const void *uniforms144, const void *limits12, uint32_t cmd_count,
               const uint32_t *words, uint32_t n_words, uint32_t W, uint32_t H, uint32_t row0,
               uint32_t nrows, uint8_t *rgba, uint8_t *cls, float *t, uint64_t *counters,
               int nthreads) {
    if (!uniforms144 || !limits12 || !rgba || W == 0 || H == 0) return RMO_E_INVALID_ARG;
    if ((uint64_t)row0 + nrows > H) return RMO_E_INVALID_ARG;
    int rc = validate_program(cmd_count, words, n_words);
    if (rc != RMO_OK) return rc;
    rmo_uniforms U;
    memcpy(&U, uniforms144, sizeof U);
    rmo_scene sc;
    sc.cmd_count = cmd_count;
    sc.words = words;
    sc.n_words = n_words;
    memcpy(&sc.lim, limits12, sizeof sc.lim);
    uint64_t n_step = 0, n_hit = 0, n_floor = 0, n_sample = 0;
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
    else omp_set_num_threads(omp_get_num_procs());
#else
    (void)nthreads;
#endif
    /* work items = 64-pixel chunks of a row, so that single rows also use every core */
    const int64_t chunks_per_row = ((int64_t)W + 63) / 64;
    const int64_t n_chunks = chunks_per_row * (int64_t)nrows;
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : n_step, n_hit, n_floor, n_sample)
    for (int64_t ch = 0; ch < n_chunks; ++ch) {
        rmo_counters cnt = {0, 0, 0, 0};
        const int64_t r = ch / chunks_per_row;
        const uint32_t x0 = (uint32_t)(ch % chunks_per_row) * 64u;
        const uint32_t x1 = x0 + 64u < W ? x0 + 64u : W;
        const uint32_t py = row0 + (uint32_t)r;
        for (uint32_t px = x0; px < x1; ++px) {
            size_t pix = (size_t)r * W + px;
            shade_pixel(&U, &sc, W, H, px, py, rgba + pix * 4, cls ? cls + pix * 16 : NULL,
                        t ? t + pix * 16 : NULL, &cnt);
        }
        n_step += cnt.n_step; n_hit += cnt.n_hit; n_floor += cnt.n_floor; n_sample += cnt.n_sample;
    }
    if (counters) {
        counters[0] = n_step; counters[1] = n_hit; counters[2] = n_floor; counters[3] = n_sample;
    }
    return RMO_OK;
}

/* map_scene at n points (xyz interleaved) — for the interpreter property tests. */
int rmo_map(const void *limits12, uint32_t cmd_count, const uint32_t *words, uint32_t n_words,
            const float *pts, uint64_t n, float *out) {
    if (!limits12 || !pts || !out) return RMO_E_INVALID_ARG;
    int rc = validate_program(cmd_count, words, n_words);
    if (rc != RMO_OK) return rc;
    rmo_scene sc;
    sc.cmd_count = cmd_count;
    sc.words = words;
    sc.n_words = n_words;
    memcpy(&sc.lim, limits12, sizeof sc.lim);
    for (uint64_t i = 0; i < n; ++i) out[i] = map_scene(&sc, pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]);
    return RMO_OK;
}

/* One ray (origin, unit direction) through ray_march — for K1/K2-style known-answer tests.
 * out = {class, r, g, b, t, n_step}. */
int rmo_trace(const void *limits12, uint32_t cmd_count, const uint32_t *words, uint32_t n_words,
              const float o[3], const float d[3], float out[6]) {
    if (!limits12 || !o || !d || !out) return RMO_E_INVALID_ARG;
    int rc = validate_program(cmd_count, words, n_words);
    if (rc != RMO_OK) return rc;
    rmo_scene sc;
    sc.cmd_count = cmd_count;
    sc.words = words;
    sc.n_words = n_words;
    memcpy(&sc.lim, limits12, sizeof sc.lim);
    rmo_counters cnt = {0, 0, 0, 0};
    float col[3], t = 0.0f;
    int c = ray_march(&sc, o, d, col, &t, &cnt);
    out[0] = (float)c; out[1] = col[0]; out[2] = col[1]; out[3] = col[2]; out[4] = t;
    out[5] = (float)cnt.n_step;
    return RMO_OK;
}

int rmo_validate(uint32_t cmd_count, const uint32_t *words, uint32_t n_words) {
    return validate_program(cmd_count, words, n_words);
}

int rmo_num_threads(void) {
#ifdef _OPENMP
    return omp_get_num_procs();
#else
    return 1;
#endif
}
