/* Oracle-B: single-threaded CPU restatement of the reference splat shaders.
 *
 * TEST INFRASTRUCTURE ONLY. Nothing in the product (gaussian_splat_b200/) may
 * include, link or call this. Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference leg use it, as the checker.
 *
 * PARITY UNPINNED: the reference ships no tests, golden vectors or fixtures
 * for this path and its Metal shaders cannot run on
 * this box. The pins that exist are (a) Oracle-A (oracle/_ref: the .metal text
 * itself compiled as C++ under a shim) and (b) the reference ply_loader.cpp
 * compiled unmodified; see oracle/Makefile.
 *
 * What is restated (file:line relative to /root/reference):
 *   S1 per splat      shaders/gaussian_splat_tile.metal:40-157
 *                     (= shaders/gaussian_splat_50layer.metal:50-146)
 *   S2 per fragment   shaders/gaussian_splat_tile.metal:178-206
 *                     shaders/gaussian_splat_50layer.metal:149-176
 *   S3 TILE           shaders/gaussian_splat_tile.metal:215-267
 *   S3 LAYER50        shaders/gaussian_splat_50layer.metal:180-223
 *   MLAB              shaders/gaussian_splat.metal:201-291 (K-buffer insert + merge), :329-361 (resolve)
 *   uniforms / VP     src/instanced_splat_renderer.mm:453-469
 *   camera            src/trackball_camera.mm:127-163
 *   SplatInstance     src/instanced_splat_renderer.h:6-11
 * arranged in the build's stages (preprocess, scan, duplicate, stable sort,
 * tile ranges, per-tile blend) so each stage's output can be compared
 * bit-for-bit with the CUDA path. Pinned decisions D1-D12: DESIGN.md.
 */
#ifndef ORACLE_B_H
#define ORACLE_B_H
#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { ORB_MODE_TILE = 0, ORB_MODE_LAYER50 = 1, ORB_MODE_MLAB = 2 };
enum { ORB_ORDER_NEAR_FIRST = 0, ORB_ORDER_REFERENCE = 1 };

typedef struct {
    int32_t mode;           /* ORB_MODE_*                                   */
    int32_t max_fragments;  /* K: 24 / 32 / 50                              */
    int32_t order;          /* ORB_ORDER_*                                  */
    int32_t sh_degree;      /* 0 = reference parity; 1..3 build extension   */
    int32_t tile_row_begin; /* band of tile rows rendered: [begin, end)     */
    int32_t tile_row_end;   /* <= 0 means "all rows"                        */
    int32_t cap_policy;     /* 0 = mode default (D4); 1 = arrival; 2 = sorted */
    int32_t tile_test;      /* 1 = D9' (ellipse-clipped tile rectangle + per-tile test; what the CUDA path does), 0 = plain quad AABB (D9) */
} orb_opts;

/* 48-byte per-splat record, identical to the CUDA path's layout. */
typedef struct {
    float    cx, cy;        /* centre, pixel units, y DOWN                  */
    float    ux, uy;        /* u = dx*ux + dy*uy                            */
    float    vx, vy;        /* v = dx*vx + dy*vy                            */
    float    opacity;       /* color.a (fp32)                               */
    uint32_t rg;            /* half r | half g << 16                        */
    uint32_t bd;            /* half b | half depth << 16                    */
    float    z_front;       /* -viewPos.z                                   */
    uint32_t rect_lo;       /* tx0 | ty0 << 16                              */
    uint32_t rect_hi;       /* tx1 | ty1 << 16 (inclusive)                  */
} orb_record;

typedef struct orb_frame orb_frame;

/* view / proj are column-major float[16] (simd_float4x4 layout). instances is
 * the 56-byte AoS SplatInstance array. sh_rest may be NULL (degree 0).
 * cam_pos is only used when sh_degree > 0. Runs every stage up to tile ranges. */
orb_frame* orb_frame_create(const float* instances, uint32_t n, const float* sh_rest,
                            const float* view, const float* proj, int w, int h,
                            const orb_opts* opts);
void orb_frame_destroy(orb_frame*);
/* Threads used by orb_frame_create's per-splat loop (default 1 = the named single-thread baseline). */
void orb_set_threads(int n);

uint32_t        orb_n(const orb_frame*);
uint64_t        orb_num_pairs(const orb_frame*);
uint32_t        orb_num_tiles(const orb_frame*);
const orb_record* orb_records(const orb_frame*);        /* [n], culled ones zeroed     */
const uint32_t* orb_tiles_touched(const orb_frame*);    /* [n]                          */
const uint32_t* orb_offsets(const orb_frame*);          /* [n] exclusive scan           */
const uint64_t* orb_keys_unsorted(const orb_frame*);    /* [D]                          */
const uint32_t* orb_vals_unsorted(const orb_frame*);    /* [D]                          */
const uint64_t* orb_keys_sorted(const orb_frame*);      /* [D]                          */
const uint32_t* orb_vals_sorted(const orb_frame*);      /* [D]                          */
const uint32_t* orb_tile_ranges(const orb_frame*);      /* [T][2] = begin,end           */

/* Blend into a w*h half4 (8 B / pixel) framebuffer, rows outside the band are
 * left untouched. n_threads <= 1: the named single-thread baseline; > 1 splits
 * tile rows over std::thread (informational). Returns (pixel,splat) tests done. */
uint64_t orb_blend(const orb_frame*, uint16_t* fb_half4, int n_threads);
/* The same for the tile rows [row_begin, row_end) of the frame's band only (full-size spot checks). */
uint64_t orb_blend_rows(const orb_frame*, uint16_t* fb_half4, int row_begin, int row_end, int n_threads);

/* Host helpers restating the camera and the uniform set-up. */
void orb_look_at(const float eye[3], const float center[3], const float up[3], float out16[16]);
void orb_perspective(float fov_deg, int vw, int vh, float z_near, float z_far, float out16[16]);
void orb_mat4_mul(const float a[16], const float b[16], float out16[16]);

/* Pinned scalar pieces, exported for unit tests. */
float    orb_exp_neg_half(float q);          /* exp(-0.5*q), pinned polynomial */
uint16_t orb_float_to_half(float f);
float    orb_half_to_float(uint16_t h);

/* Stable LSD reference for the stand-alone sort entry. */
void orb_sort_pairs(uint64_t* keys, uint32_t* vals, uint64_t n);

#ifdef __cplusplus
}
#endif
#endif
