/*
 * rt_oracle.c -- CPU restatement of the reference hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library, and only as the checker or the timed CPU baseline; the
 * product (python_ray_tracer_b200 + librt_b200.so) never links, imports or calls it.
 *
 * What it restates: NumpyRenderer.raytrace_scene and everything it calls in
 * tristanGIANDO/python_ray_tracer, one scalar double-precision evaluation per ray,
 * in the reference's operation order, recursion included.  Build with
 * -ffp-contract=off (oracle/Makefile) so no multiply-add is fused: the reference's
 * NumPy ufuncs round after every operation.
 *
 * Pinning: checked against the unmodified reference imported from /root/reference
 * (tests/golden/make_golden.py wrote the fixtures in tests/golden/, including the
 * main.py scene whose 8-bit image equals the reference's committed render.png).
 *
 * Paths cited below are relative to the reference tree.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/rt_abi.h"

#ifdef _OPENMP
#include <omp.h>
#endif

typedef struct v3 {
    double x, y, z;
} v3;

typedef struct orc_scene {
    const double* f; /* sphere SoA, field-major (rt_abi.h) */
    int n;
    const double* lights; /* n_point x 3 */
    const rt_env* env;
} orc_scene;

typedef struct orc_counters {
    uint64_t traced, hits, deepest, tests;
} orc_counters;

#define FIELD(sc, fld, s) ((sc)->f[(size_t)(fld) * (size_t)(sc)->n + (size_t)(s)])

/* NumpyVector3D.dot (base.py:34-35): products summed left to right. */
static inline double dot3(v3 a, v3 b) { return (a.x * b.x) + (a.y * b.y) + (a.z * b.z); }
static inline v3 sub3(v3 a, v3 b) { return (v3){a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline v3 add3(v3 a, v3 b) { return (v3){a.x + b.x, a.y + b.y, a.z + b.z}; }
static inline v3 scale3(v3 a, double k) { return (v3){a.x * k, a.y * k, a.z * k}; }

/* NumpyVector3D.norm (base.py:61-64): multiply by the reciprocal of the length,
 * a zero length is replaced by 1. */
static inline v3 norm3(v3 a) {
    double mag = sqrt(dot3(a, a));
    double inv = 1.0 / (mag == 0.0 ? 1.0 : mag);
    return scale3(a, inv);
}

static inline double clip01(double v) { return v < 0.0 ? 0.0 : (v > 1.0 ? 1.0 : v); }

/* NumpySphere.intersect (shape.py:28-51). */
static double sphere_distance(v3 o, v3 d, v3 c, double r) {
    double b = 2.0 * dot3(d, sub3(o, c));
    double cc = dot3(c, c) + dot3(o, o) - 2.0 * dot3(c, o) - (r * r);
    double disc = (b * b) - (4.0 * cc);
    double sq = sqrt(disc > 0.0 ? disc : 0.0);
    double h0 = (-b - sq) / 2.0;
    double h1 = (-b + sq) / 2.0;
    double h = (h0 > 0.0 && h0 < h1) ? h0 : h1;
    return (disc > 0.0 && h > 0.0) ? h : RT_FARAWAY;
}

static inline v3 sphere_center(const orc_scene* sc, int s) {
    return (v3){FIELD(sc, RT_F_CX, s), FIELD(sc, RT_F_CY, s), FIELD(sc, RT_F_CZ, s)};
}

/* Texture.get_color / TextureChecker.get_color (shader.py:13-32).  astype(int)
 * truncates toward zero; Python's % 2 maps negatives into {0,1}. */
static v3 texture_color(const orc_scene* sc, int s, v3 m) {
    if (FIELD(sc, RT_F_TEXTURE, s) != 0.0) {
        long long ix = (long long)(m.x * 2.0);
        long long iz = (long long)(m.z * 2.0);
        double on = ((ix & 1LL) == (iz & 1LL)) ? 1.0 : 0.0;
        return (v3){on, on, on};
    }
    return (v3){FIELD(sc, RT_F_COLOR_R, s), FIELD(sc, RT_F_COLOR_G, s),
                FIELD(sc, RT_F_COLOR_B, s)};
}

/* NumpyShader._calculate_physical_specular (shader.py:246-320); grey value. */
static double physical_specular(v3 n, v3 to_light, v3 to_eye, double rough, double gain) {
    const double eps = 1e-8;
    v3 l = norm3(to_light);
    v3 v = norm3(to_eye);
    v3 h = norm3(add3(l, v));
    double ndv = clip01(dot3(n, v));
    double ndh = clip01(dot3(n, h));
    double vdh = clip01(dot3(v, h));
    double ndl = clip01(dot3(n, l));

    double f0 = pow((1.5 - 1.0) / (1.5 + 1.0), 2.0);
    double fres = f0 + (1.0 - f0) * pow(1.0 - vdh, 5.0);

    double alpha = pow(rough, 2.0);
    double a2 = pow(alpha, 2.0);
    double den = (ndh * ndh) * (a2 - 1.0) + 1.0;
    double dist = a2 / (M_PI * ((den * den) + eps));

    double ql = clip01(dot3(n, l));
    double g1l = 2.0 * ql / (ql + sqrt(a2 + (1.0 - a2) * (ql * ql)) + eps);
    double qv = clip01(dot3(n, v));
    double g1v = 2.0 * qv / (qv + sqrt(a2 + (1.0 - a2) * (qv * qv)) + eps);
    double geo = g1l * g1v;

    double base = (fres * dist * geo) / (4.0 * ndv + eps);
    double glint = pow(1.0 - ndv, 2.5);
    glint = glint * ndl;
    double total = base + gain * glint;
    return ndv <= 0.0 ? 0.0 : total;
}

/* NumpyShader._calculate_physical_iridescence (shader.py:186-232). */
static v3 iridescence(v3 n, v3 to_eye, double gain) {
    const double weight = 0.1, thick = 0.3, film_ior = 1.4;
    double va = clip01(dot3(n, to_eye));
    double af = fabs(va - 0.5) * 2.0;
    double phase = af * M_PI * thick * 10.0;
    double ip = sin(phase);
    double hs = (film_ior - 1.0) / 2.0;
    double r = ip * hs + (1.0 - hs) * (1.0 - ip);
    double g = ip * (1.0 - hs) + hs * (1.0 - ip);
    double b = 0.5 + 0.5 * ip;
    return (v3){r * weight * gain, g * weight * gain, b * weight * gain};
}

static v3 trace(const orc_scene* sc, v3 o, v3 d, int generation, orc_counters* k);

/* NumpyShader._calculate_shadow (shader.py:114-128) toward one light: the point is
 * lit when its own sphere's distance along the shadow ray equals the minimum over all
 * shapes (so "nothing hit at all" is lit, and there is no test against the distance
 * to the light). */
static int in_light(const orc_scene* sc, int s, v3 nudged, v3 to_light, orc_counters* k) {
    double own = RT_FARAWAY, nearest = RT_FARAWAY;
    for (int j = 0; j < sc->n; ++j) {
        double u = sphere_distance(nudged, to_light, sphere_center(sc, j), FIELD(sc, RT_F_RADIUS, j));
        if (j == s) own = u;
        if (j == 0 || u < nearest) nearest = u;
    }
    k->tests += (uint64_t)sc->n;
    return own == nearest;
}

/* NumpyShader.create (shader.py:63-112) for one hit of sphere s at distance dist. */
static v3 shade_hit(const orc_scene* sc, int s, v3 o, v3 d, double dist, int generation,
                    orc_counters* k) {
    const rt_env* env = sc->env;
    v3 c = sphere_center(sc, s);
    double radius = FIELD(sc, RT_F_RADIUS, s);
    double diffuse_gain = FIELD(sc, RT_F_DIFFUSE_GAIN, s);
    double specular_gain = FIELD(sc, RT_F_SPECULAR_GAIN, s);
    double rough = FIELD(sc, RT_F_SPECULAR_ROUGHNESS, s);
    double irid_gain = FIELD(sc, RT_F_IRIDESCENCE_GAIN, s);
    v3 eye = {env->eye[0], env->eye[1], env->eye[2]};

    v3 m = add3(o, scale3(d, dist));                    /* :73 */
    v3 n = scale3(sub3(m, c), 1.0 / radius);            /* :74 */
    v3 to_eye = norm3(sub3(eye, m));                    /* :76 */
    v3 nudged = add3(m, scale3(n, 0.0001));             /* :77 */

    int n_lights = (env->flags & RT_FLAG_ALL_POINT_LIGHTS) ? env->n_point_lights : 1;
    v3 tex = texture_color(sc, s, m);

    v3 color = {0.004, 0.004, 0.004};                   /* :86, :183-184 */
    double spec_lit_sum = 0.0;
    int lit0 = 0;
    for (int li = 0; li < n_lights; ++li) {
        v3 lp = {sc->lights[3 * li], sc->lights[3 * li + 1], sc->lights[3 * li + 2]};
        v3 to_light = norm3(sub3(lp, m));               /* :75 */
        int lit = in_light(sc, s, nudged, to_light, k); /* :79 */
        double litf = lit ? 1.0 : 0.0;
        if (li == 0) lit0 = lit;
        /* :88, :130-141 */
        double lambert = dot3(n, to_light);
        lambert = lambert > 0.0 ? lambert : 0.0;
        color.x += tex.x * lambert * litf * diffuse_gain;
        color.y += tex.y * lambert * litf * diffuse_gain;
        color.z += tex.z * lambert * litf * diffuse_gain;
        double spec = physical_specular(n, to_light, to_eye, rough, specular_gain); /* :100 */
        if (n_lights == 1) {
            spec_lit_sum = spec; /* gated below together with the reflection, as :106 does */
        } else {
            spec_lit_sum += spec * litf;
        }
    }

    /* :90, :143-161 -- mirror direction, then recurse from the nudged point. */
    v3 rd = norm3(sub3(d, scale3(scale3(n, 2.0), dot3(d, n))));
    v3 refl = trace(sc, nudged, rd, generation + 1, k);

    /* :98, :234-244 */
    double ny = (n.x * 0.0) + (n.y * 1.0) + (n.z * 0.0);
    ny = ny > 0.0 ? ny : 0.0;
    double dome = 0.0;
    for (int li = 0; li < env->n_dome_lights; ++li) dome += env->dome_intensity[li] * ny;
    color.x += env->dome_color[0] * dome;
    color.y += env->dome_color[1] * dome;
    color.z += env->dome_color[2] * dome;

    /* :106 */
    double lit0f = lit0 ? 1.0 : 0.0;
    if (n_lights == 1) {
        color.x += (spec_lit_sum + refl.x * 0.5) * specular_gain * lit0f;
        color.y += (spec_lit_sum + refl.y * 0.5) * specular_gain * lit0f;
        color.z += (spec_lit_sum + refl.z * 0.5) * specular_gain * lit0f;
    } else {
        /* extension: every light's specular gated by its own shadow test, the
         * reflection gated by light 0 as in the reference. */
        color.x += (spec_lit_sum + refl.x * 0.5 * lit0f) * specular_gain;
        color.y += (spec_lit_sum + refl.y * 0.5 * lit0f) * specular_gain;
        color.z += (spec_lit_sum + refl.z * 0.5 * lit0f) * specular_gain;
    }

    v3 film = iridescence(n, to_eye, irid_gain);        /* :110 */
    return add3(color, film);
}

/* NumpyRenderer.raytrace_scene (base.py:91-121) for one ray.  Every shape whose
 * distance equals the minimum contributes (exact ties add up, as the masks do). */
static v3 trace(const orc_scene* sc, v3 o, v3 d, int generation, orc_counters* k) {
    v3 black = {0.0, 0.0, 0.0};
    if (sc->env->max_depth > 0 && generation >= sc->env->max_depth) return black;
    k->traced += 1;
    k->tests += (uint64_t)sc->n;
    if ((uint64_t)(generation + 1) > k->deepest) k->deepest = (uint64_t)(generation + 1);

    double nearest = RT_FARAWAY;
    for (int s = 0; s < sc->n; ++s) {
        double t = sphere_distance(o, d, sphere_center(sc, s), FIELD(sc, RT_F_RADIUS, s));
        if (s == 0 || t < nearest) nearest = t;
    }
    if (nearest == RT_FARAWAY) return black;

    v3 color = black;
    for (int s = 0; s < sc->n; ++s) {
        double t = sphere_distance(o, d, sphere_center(sc, s), FIELD(sc, RT_F_RADIUS, s));
        if (t == nearest) {
            k->hits += 1;
            color = add3(color, shade_hit(sc, s, o, d, t, generation, k));
        }
    }
    return color;
}

/* ------------------------------------------------------------------------------------
 * Exported checker entry points (all host memory, doubles).
 * ---------------------------------------------------------------------------------- */

/* np.linspace(start, stop, num)[i] as NumPy evaluates it: i*step + start with
 * step = (stop-start)/(num-1), the last sample overwritten with stop; num == 1
 * gives start. */
static double linspace_at(double start, double stop, int num, int i) {
    if (num <= 1) return 0.0 * (stop - start) + start;
    if (i == num - 1) return stop;
    double step = (stop - start) / (double)(num - 1);
    if (step == 0.0) return ((double)i / (double)(num - 1)) * (stop - start) + start;
    return (double)i * step + start;
}

/* NumpyRenderer.get_ray_directions (base.py:123-141) for pixel p = row*W + col. */
static v3 primary_direction(const rt_camera* cam, int64_t p) {
    int w = cam->width, h = cam->height;
    int row = (int)(p / w), col = (int)(p % w);
    double aspect = (double)w / (double)h;
    double x = linspace_at(-1.0, 1.0, w, col);
    double y = linspace_at(1.0 / aspect + 0.25, -1.0 / aspect + 0.25, h, row);
    v3 q = {x - cam->position[0], y - cam->position[1], 0.0 - cam->position[2]};
    return norm3(q);
}

double orc_sphere_distance(const double* o, const double* d, const double* c, double r) {
    return sphere_distance((v3){o[0], o[1], o[2]}, (v3){d[0], d[1], d[2]},
                           (v3){c[0], c[1], c[2]}, r);
}

/* pixels == NULL: pixel indices first .. first+n-1; out is 3 planes of n doubles. */
int orc_ray_directions(const rt_camera* cam, const int64_t* pixels, int64_t first, int64_t n,
                       double* out) {
    if (!cam || !out || n < 0 || cam->width <= 0 || cam->height <= 0) return RT_ERR_BAD_ARG;
    for (int64_t i = 0; i < n; ++i) {
        v3 d = primary_direction(cam, pixels ? pixels[i] : first + i);
        out[i] = d.x;
        out[n + i] = d.y;
        out[2 * n + i] = d.z;
    }
    return RT_OK;
}

static void merge_counters(uint64_t* stats, const orc_counters* k) {
    if (!stats) return;
#ifdef _OPENMP
#pragma omp critical
#endif
    {
        stats[0] += k->traced;
        stats[1] += k->hits;
        if (k->deepest > stats[2]) stats[2] = k->deepest;
        stats[3] += k->tests;
    }
}

/* raytrace_scene over an arbitrary ray batch.  origins == NULL broadcasts origin0.
 * threads <= 0: all cores OpenMP gives us. */
int orc_trace_rays(const double* spheres, int32_t n_spheres, const double* point_lights,
                   const rt_env* env, const double* origins, const double* origin0,
                   const double* directions, int64_t n, double* rgb, uint64_t* stats,
                   int threads) {
    if (!spheres || n_spheres <= 0 || !point_lights || !env || !directions || !rgb || n < 0)
        return RT_ERR_BAD_ARG;
    if (!origins && !origin0) return RT_ERR_BAD_ARG;
    orc_scene sc = {spheres, n_spheres, point_lights, env};
    if (stats) memset(stats, 0, RT_STATS_WORDS * sizeof(uint64_t));
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#pragma omp parallel
#endif
    {
        orc_counters k = {0, 0, 0, 0};
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 256)
#endif
        for (int64_t i = 0; i < n; ++i) {
            v3 o = origins ? (v3){origins[i], origins[n + i], origins[2 * n + i]}
                           : (v3){origin0[0], origin0[1], origin0[2]};
            v3 d = {directions[i], directions[n + i], directions[2 * n + i]};
            v3 c = trace(&sc, o, d, 0, &k);
            rgb[i] = c.x;
            rgb[n + i] = c.y;
            rgb[2 * n + i] = c.z;
        }
        merge_counters(stats, &k);
    }
    return RT_OK;
}

/* get_ray_directions + raytrace_scene for camera pixels (pixels == NULL: first..). */
int orc_render_pixels(const double* spheres, int32_t n_spheres, const double* point_lights,
                      const rt_env* env, const rt_camera* cam, const int64_t* pixels,
                      int64_t first, int64_t n, double* rgb, uint64_t* stats, int threads) {
    if (!spheres || n_spheres <= 0 || !point_lights || !env || !cam || !rgb || n < 0)
        return RT_ERR_BAD_ARG;
    if (cam->width <= 0 || cam->height <= 0) return RT_ERR_BAD_ARG;
    orc_scene sc = {spheres, n_spheres, point_lights, env};
    if (stats) memset(stats, 0, RT_STATS_WORDS * sizeof(uint64_t));
    v3 o = {cam->position[0], cam->position[1], cam->position[2]};
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#pragma omp parallel
#endif
    {
        orc_counters k = {0, 0, 0, 0};
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 256)
#endif
        for (int64_t i = 0; i < n; ++i) {
            v3 d = primary_direction(cam, pixels ? pixels[i] : first + i);
            v3 c = trace(&sc, o, d, 0, &k);
            rgb[i] = c.x;
            rgb[n + i] = c.y;
            rgb[2 * n + i] = c.z;
        }
        merge_counters(stats, &k);
    }
    return RT_OK;
}

/* NumpyRenderer.save_image arithmetic (base.py:143-151): truncating quantisation,
 * planar doubles -> interleaved RGB bytes. */
int orc_quantize(const double* rgb, int64_t n, uint8_t* out) {
    if (!rgb || !out || n < 0) return RT_ERR_BAD_ARG;
    for (int64_t i = 0; i < n; ++i)
        for (int c = 0; c < 3; ++c) out[3 * i + c] = (uint8_t)(255.0 * clip01(rgb[c * n + i]));
    return RT_OK;
}

int orc_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
