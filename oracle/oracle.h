/*
 * oracle.h -- C interface of the CPU oracle.
 *
 * TEST INFRASTRUCTURE ONLY.  The oracle is a scalar CPU restatement of the reference's per-ray
 * algorithm (shaders/tracer.cu + shaders/tracer.cuh of Ray-Studio2/gaussian-ray-tracing).  It may be
 * loaded only by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
 * The product (libgrt.so) never links, loads or calls it.
 *
 * PARITY UNPINNED: the reference ships no golden vectors, tests or CPU path for this hot path
 *  and its OptiX device code cannot be built here.  The pins we do have:
 *   - the reference's own src/GaussianData.cpp and src/Camera.cpp compiled unchanged into
 *     oracle/_ref/libref_host.so (PLY parse + activation, U/V/W camera frame) -- see ref_shim.cpp;
 *   - analytic known-answer tests derived from the formulas in tracer.cuh (tests/test_oracle_kat.py).
 */
#ifndef GRT_ORACLE_H
#define GRT_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Same 236-byte record as the reference's GaussianParticle (src/GaussianData.h:12-20). */
typedef struct OrcParticle {
    float position[3];
    float scale[3];    /* already exp()'d */
    float rotation[4]; /* memory order x,y,z,w (glm::quat) */
    float opacity;     /* already sigmoid()'d */
    float sh[48];      /* 16 x float3, coefficient-major */
} OrcParticle;

/* The numeric fields of the reference's Params (src/Parameters.h:42-74). */
typedef struct OrcParams {
    uint32_t width, height, sh_degree_max;
    float eye[3], U[3], V[3], W[3];
    float t_min, t_max, minTransmittance, alpha_min;
    int32_t mode_fisheye;
    int32_t type; /* MeshType: 0 MIRROR, 1 NORMAL, 2 GLASS */
} OrcParams;

/* One reflection primitive, already prepared like the reference host does
 * (world-space vertices = transform * v; normals = mat3(transform) * n, src/GaussianTracer.cpp:653-709). */
typedef struct OrcMesh {
    const float* vertices;  /* n_vertices x 3, world space */
    const float* normals;   /* n_vertices x 3 */
    const uint32_t* faces;  /* n_faces x 3 */
    uint32_t n_vertices, n_faces;
} OrcMesh;

typedef struct OrcStats {
    uint64_t rays;          /* pixels rendered */
    uint64_t segments;      /* traceGaussians calls */
    uint64_t rounds;        /* traceGPs calls (k-buffer gathers) */
    uint64_t hits_gathered; /* h_g: k-buffer entries consumed by the integration loop (tracer.cuh:353) */
    uint64_t hits_shaded;   /* h_s: entries with alpha_min < alpha (tracer.cuh:361) */
    uint64_t flagged;       /* pixels flagged tie/threshold-ambiguous */
    uint64_t nodes_visited; /* CPU BVH statistics (not part of the reference) */
    uint64_t proxies_tested;
    double seconds;         /* wall time of the frame loop */
    uint32_t threads;
} OrcStats;

typedef struct OrcScene OrcScene;

/* flags in the per-pixel flag image */
#define ORC_FLAG_TIE 1u        /* two gathered face hits within 4 ulp, or a hit within 4 ulp of a segment bound */
#define ORC_FLAG_ALPHA 2u      /* |alpha - alpha_min| <= 5e-7 for some hit */
#define ORC_FLAG_TRANSMIT 4u   /* |T - minTransmittance| <= 5e-8 at some check */
#define ORC_FLAG_MESH 8u       /* mesh closest hit ambiguous (two triangles within 4 ulp) */

/* Scene = particles (+ proxy transform of GaussianTracer.cpp:297-317 with alpha_min_build) + CPU BVH. */
OrcScene* orc_scene_create(const OrcParticle* particles, uint64_t n, float alpha_min_build);
void orc_scene_destroy(OrcScene* s);
/* Replace the reflection primitives (0 meshes == null mesh handle: every mesh ray misses). */
void orc_scene_set_meshes(OrcScene* s, const OrcMesh* meshes, uint32_t n_meshes);
uint64_t orc_scene_live_count(const OrcScene* s);

typedef struct OrcRenderArgs {
    OrcParams params;
    uint32_t K;             /* k-buffer size; reference = 7 (MaxNumHitPerTrace) */
    int32_t brute_force;    /* 1: test every particle / triangle, no BVH */
    int32_t use_triangles;  /* 1: intersect the 20 proxy triangles (Moller-Trumbore) instead of 10 slabs */
    uint32_t threads;       /* 0 = hardware_concurrency */
    uint32_t x0, y0, w, h;  /* sub-rectangle to render (w==0: full frame) */
    uint8_t* out_rgb8;      /* width*height*3 or NULL; indexed y*width+x like the reference */
    float* out_rgb32f;      /* width*height*3 or NULL: accumColor before clamp/quantise */
    uint8_t* out_flags;     /* width*height or NULL */
    uint16_t* out_hg;       /* width*height or NULL: gathered hits per pixel (saturating) */
    /* hit-ID traces: for pixel trace_pixels[i] (y*width+x) the first trace_max consumed particle IDs */
    const uint32_t* trace_pixels;
    uint32_t n_trace_pixels, trace_max;
    uint32_t* trace_ids;    /* n_trace_pixels * trace_max */
    uint32_t* trace_counts; /* n_trace_pixels: total consumed (may exceed trace_max) */
} OrcRenderArgs;

int orc_render(const OrcScene* s, const OrcRenderArgs* args, OrcStats* stats);

/* --- known-answer hooks (each restates one reference function) --- */
float orc_compute_response(const OrcParticle* p, const float o[3], const float d[3]);          /* tracer.cuh:187-214 */
void orc_sh_radiance(const OrcParticle* p, const float d[3], uint32_t deg, float out[3]);      /* tracer.cuh:216-264 */
uint8_t orc_quantize(float x);                                                                 /* tracer.cuh:68-73 */
int orc_get_ray(const OrcParams* prm, uint32_t x, uint32_t y, float o[3], float d[3]);         /* tracer.cuh:115-165; returns 0 if no ray */
/* proxy intersection of one particle: returns number of face hits (0 or 2) and their t */
int orc_proxy_hit(const OrcParticle* p, float alpha_min_build, const float o[3], const float d[3],
                  int use_triangles, float t_out[2]);
void orc_icosahedron(float verts[36], uint32_t idx[60]);                                       /* Icosahedron.h:11-38 */
void orc_refract(const float d[3], const float n[3], float eta, float out_dir[3], float* t_hit, uint32_t* bounces); /* tracer.cuh:432-465 */

/* mesh preparation restated from src/geometry/Primitives.cpp and src/GaussianTracer.cpp:653-709 */
void orc_transform_mesh(const float* verts, const float* normals, uint32_t n, const float transform_colmajor[16],
                        float* verts_world, float* normals_out);

/* batch forms of the hooks above (tests/test_ref_pins.py: oracle vs reference-compiled code on 1e5 random inputs) */
void orc_compute_response_batch(const OrcParticle* ps, uint64_t n, const uint32_t* pidx, const float* o, const float* d, float* out);
void orc_sh_radiance_batch(const OrcParticle* ps, uint64_t n, const uint32_t* pidx, const float* d, uint32_t deg, int normalise, float* out);
void orc_quantize_batch(uint64_t n, const float* x, uint8_t* out);
void orc_get_rays(const OrcParams* prm, uint64_t n, const uint32_t* xs, const uint32_t* ys, float* o, float* d, uint8_t* valid);
void orc_refract_batch(uint64_t n, const float* d, const float* nrm, float eta, float* out_dir, float* t_hit_io, uint32_t* bounces_io);
void orc_mirror_batch(uint64_t n, const float* d, const float* nrm, float* out_dir, uint32_t* bounces_io);
void orc_proxy_data(const OrcParticle* ps, uint64_t n, float alpha_min_build, float* A9, float* s_out, float* wv36, float* box6, uint8_t* live);
void orc_proxy_hit_batch(const OrcParticle* ps, uint64_t n, const uint32_t* pidx, float alpha_min_build, const float* o, const float* d,
                         float* t2, uint8_t* hit);
void orc_soup_hits(uint64_t n, const float* verts36, const uint32_t* idx60, const float* o, const float* d, float* t2, uint8_t* hit);

#ifdef __cplusplus
}
#endif
#endif
