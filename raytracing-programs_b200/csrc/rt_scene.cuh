// csrc/rt_scene.cuh -- device-resident scene: type-sorted structure of arrays that
// replaces the reference's linked list / shape_t array (C/objects.h:15-48,
// C/raycast.c:87-102).  Geometry is split from materials so the intersection loops
// touch 16 B per sphere; the original file-order index travels with every primitive
// because the nearest-hit tie-break and the self-skip are defined on it
// (C/raycast.c:450,456).
#pragma once
#include "rt_math.cuh"
#include "rt_qindex.cuh"

#define RT_MAX_DEPTH_CAP 64
struct DevScene {
  int n_shapes, n_spheres, n_planes, n_quadrics, n_lights;
  int max_depth;            // MAX_RECURSION of C/raycast.c:14 (7); rt_set_max_depth changes it (dormant deep-bounce mode)
  float cam_w, cam_h;
  const float4 *sph;        // [n_spheres] {cx, cy, cz, r}
  const float4 *sph_hot;    // [n_spheres padded to 4] reject-filter operands {cx, cy, cz, fl(r*r)}; pads never pass
  const float4 *sph_prim;   // [n_spheres] primary-ray constants {0-cx, 0-cy, 0-cz, c} (origin is exactly 0)
  int n_csph;               // sphere slots 0 .. n_csph-1 are (also) in c_sph / c_prim; padded to a multiple of 8 there
  const int *sph_idx;       // [n_spheres] original index
  const float4 *pln;        // [2*n_planes] {px,py,pz,0} {nx,ny,nz,0}
  const int *pln_idx;
  const float4 *quad;       // [3*n_quadrics] {a,b,c,d} {e,f,g,h} {i,j,0,0}
  const float4 *quad_hot;   // [3*(n_quadrics+1)] reject-filter operands {A,B,C,D} {E,F,G,H} {I,J,S2,S1}
  const float4 *quad_ball;  // [n_quadrics padded to 32 + 4] bounding ball {cx, cy, cz, a} of each quadric slot for the pre-test of the quadric
                            // scan (rt_kernels.cuh, "Quadrics"): radius a + quad_ball_eps |P - c|; a = +inf for a quadric without a
                            // certified ball (always passes), -inf for pads (never pass)
  const float4 *quad_ball_prim;   // [n_quadrics] the same test for the primary rays (origin exactly 0): {cx, cy, cz, -kappa}
  float quad_ball_eps;      // M: the largest miss_rel among the certified ellipsoids (rt_qindex.cuh), rounded up
  const float4 *quad_prim;  // [3*n_quadrics] filter constants for the origin (0,0,0): {As,Bs,Cs,Ds} {Es,Fs,m0,eb} {gx,gy,gz,0}
  const int *quad_idx;
  const float4 *mat;        // [2*n_shapes] by original index: {Kd.rgb, reflectivity} {Ks.rgb, opacity}
  const int2 *slot;         // [n_shapes] by original index: {type, slot in its type array}
  const float4 *lights;     // [4*n_lights] {pos, type} {color, cos_theta} {radial a0 a1 a2, angular a0} {dir, cull flag}
  // Light buffer (opt-in exact acceleration of the sphere part of shadow queries):
  // per light and distance level a cube map of RT_LB_RES^2 cells per face; a cell lists every sphere the
  // reference could report as hit by a shadow ray whose direction seen from the light falls in the cell.
  const unsigned *lb_start; // [n_lights * RT_LB_LEVELS * 6 * RT_LB_RES^2 + 1] CSR starts; NULL = no light buffer
  const int *lb_items;      // sphere slots
  const float4 *lb_hot;     // sph_hot of each item, so walking a list costs no dependent load
  const float *lb_caps;     // [n_lights * RT_LB_LEVELS] squared |P - L| limit of each level, ascending (-1: unusable)
  // Bounding-sphere hierarchy over the spheres (same opt-in mode): nearest hit of reflection rays and the
  // fallback of shadow rays.  Node k: bvh_node[k] = {centre, radius of a sphere holding all member spheres},
  // bvh_kids[k] = {left, right} child nodes, or for a leaf {-(first + 1), count}: sphere slots first .. first+count-1
  // (in this mode the host lays the spheres out in leaf order).
  const float4 *bvh_node;
  const int2 *bvh_kids;
  const float4 *bvh_pair;   // [2k], [2k+1]: bvh_node of the left and right child of inner node k (one fetch per traversal step)
  // The same kind of hierarchy over the ellipsoids among the quadrics (rt_qindex.cuh): they take quadric slots
  // 0 .. n_qidx-1 in leaf order, every other quadric follows and is scanned as before.  qbvh_node = NULL or
  // n_qidx = 0: no quadric index.
  const float4 *qbvh_node;
  const int2 *qbvh_kids;
  const float4 *qbvh_marg;  // per node {miss_rel, behind_abs, t_rel, t_abs}: the largest margins its members need
  const float4 *qbvh_pair, *qbvh_mpair;   // children's node and margin entries stored with the parent, as bvh_pair
  int n_qidx;
  float q_behind_rel;
  // Light buffer over the indexed ellipsoids (same layout as lb_*; items are quadric slots); NULL: shadow rays walk
  // the hierarchy instead.
  const unsigned *lbq_start;
  const int *lbq_items;
  const float *lbq_caps;
  // The same cube map with the camera (0,0,0) as apex, for the primary rays: level 0 of "light" 0 (lb_* layout).
  const unsigned *cam_start;   // spheres; NULL: primary rays walk the hierarchies
  const int *cam_items;
  const unsigned *camq_start;  // indexed ellipsoids; NULL: hierarchy
  const int *camq_items;
};
#define RT_LB_RES 32
#define RT_LB_LEVELS 6

// Sphere reject-filter operands in __constant__ memory (the scene "staged into constant memory", north_star item 2).
// Why constant memory and not L1-cached global loads: every thread of a warp scans the same sphere at the same time, so
// an operand read from the constant bank arrives in a UNIFORM register, and an FFMA with a uniform operand reads two
// register-file operands instead of three.  On B200 a three-register FFMA issues at 0.61 per clock and scheduler, the
// uniform-operand form at 0.96 (tools/scan_shapes.cu, profiles/r02_scan_shapes.md) -- the round-1 scan was capped by
// exactly that.  Slot order = sph_hot order.  Entry: {cx, cy, cz, rho}, rho = fl(r^2) (1 + 2^-18) + 2^-19 |C|^2 rounded
// up (host, double); pads hold rho = -inf and never pass.  Scenes with more spheres than the bank holds scan the
// remainder from global memory with the round-1 loop.
#define RT_CSPH_CAP 2560
__constant__ float4 c_sph[RT_CSPH_CAP];
__constant__ float c_prim[RT_CSPH_CAP];   // primary rays (origin exactly 0): the filter constant of sph_prim[].w

// Cube-map cell of direction u (any length): face = 2 * major axis + (negative ? 1 : 0), the other two
// components divided by |major| map [-1, 1] onto RT_LB_RES cells.  Returns -1 for a zero or NaN vector.
// The host builds the cells with the same mapping in double and a margin that swallows float rounding.
__host__ __device__ inline int lb_cell(float ux, float uy, float uz) {
  const float ax = fabsf(ux), ay = fabsf(uy), az = fabsf(uz);
  float m, s, t;
  int face;
  if (ax >= ay && ax >= az) { m = ax; s = uy; t = uz; face = ux < 0.0f ? 1 : 0; }
  else if (ay >= az) { m = ay; s = ux; t = uz; face = uy < 0.0f ? 3 : 2; }
  else { m = az; s = ux; t = uy; face = uz < 0.0f ? 5 : 4; }
  if (!(m > 0.0f) || !(m < 3.0e38f)) return -1;
#ifdef __CUDA_ARCH__
  // one approximate reciprocal instead of two IEEE divisions: its 2 ulp are nothing next to the 3e-4 rad the host adds to every cell
  const float inv = __fdividef(1.0f, m);
  int i = (int)floorf((s * inv + 1.0f) * (0.5f * RT_LB_RES)), j = (int)floorf((t * inv + 1.0f) * (0.5f * RT_LB_RES));
#else
  int i = (int)floorf((s / m + 1.0f) * (0.5f * RT_LB_RES)), j = (int)floorf((t / m + 1.0f) * (0.5f * RT_LB_RES));
#endif
  i = i < 0 ? 0 : i > RT_LB_RES - 1 ? RT_LB_RES - 1 : i;
  j = j < 0 ? 0 : j > RT_LB_RES - 1 ? RT_LB_RES - 1 : j;
  return (face * RT_LB_RES + j) * RT_LB_RES + i;
}

// Hit-point queue (one wavefront level): 48 B per entry, structure of float4 arrays, plus one
// index list per work class.  The producer files every entry under the class of its estimated
// number of live rays (reflection ray + lights in front of the surface); the consumer walks the
// lists class by class, so the 32 hit points a warp shades together need about the same number
// of ray slots.  The class is only a scheduling hint: the consumer sizes its scan from the
// exact count it computes itself.
#ifndef RT_NCLS
#define RT_NCLS 8     /* one class per scan variant of the widest kernel (17, 15, ... 3 ray slots) */
#endif
struct HitQueue {
  float4 *r0;   // P.xyz, pixel (tile-local index, int bits)
  float4 *r1;   // N.xyz, object (original index, int bits)
  float4 *r2;   // V.xyz (direction of the ray that made the hit), throughput
  int *cls;     // [RT_NCLS][cap] entry numbers per class
  int *ctl;     // [0] entries, [1..RT_NCLS] entries per class, [RT_CTL_CURSOR] work cursor of the consuming launch
  int cap;
};
#define RT_CTL_CURSOR (1 + RT_NCLS)
#define RT_CTL_STRIDE 16
static_assert(RT_CTL_CURSOR < RT_CTL_STRIDE, "queue control block too small for RT_NCLS classes + the work cursor");

// Device counters (unsigned long long each).
enum {
  CT_RAYS_REFLECTION = 0, CT_RAYS_SHADOW, CT_HIT_POINTS, CT_PRIMARY_HITS,
  CT_SPH_O, CT_SPH_D, CT_PLN_O, CT_PLN_D, CT_QUAD_O, CT_QUAD_D,
  CT_LIGHT_SAMPLES, CT_CANDIDATES, CT_SLOTS, CT_BALL_O, CT_BALL_D, CT_COUNT
};
