// csrc/rt_scene.cuh -- device-resident scene: type-sorted structure of arrays that
// replaces the reference's linked list / shape_t array (C/objects.h:15-48,
// C/raycast.c:87-102).  Geometry is split from materials so the intersection loops
// touch 16 B per sphere; the original file-order index travels with every primitive
// because the nearest-hit tie-break and the self-skip are defined on it
// (C/raycast.c:450,456).
#pragma once
#include "rt_math.cuh"

struct DevScene {
  int n_shapes, n_spheres, n_planes, n_quadrics, n_lights;
  float cam_w, cam_h;
  const float4 *sph;        // [n_spheres] {cx, cy, cz, r}
  const float4 *sph_hot;    // [n_spheres padded to 4] reject-filter operands {cx, cy, cz, fl(r*r)}; pads never pass
  const float4 *sph_prim;   // [n_spheres] primary-ray constants {0-cx, 0-cy, 0-cz, c} (origin is exactly 0)
  const int *sph_idx;       // [n_spheres] original index
  const float4 *pln;        // [2*n_planes] {px,py,pz,0} {nx,ny,nz,0}
  const int *pln_idx;
  const float4 *quad;       // [3*n_quadrics] {a,b,c,d} {e,f,g,h} {i,j,0,0}
  const float4 *quad_hot;   // [3*(n_quadrics+1)] reject-filter operands {A,B,C,D} {E,F,G,H} {I,J,S2,S1}
  const float4 *quad_prim;  // [3*n_quadrics] filter constants for the origin (0,0,0): {As,Bs,Cs,Ds} {Es,Fs,m0,eb} {gx,gy,gz,0}
  const int *quad_idx;
  const float4 *mat;        // [2*n_shapes] by original index: {Kd.rgb, reflectivity} {Ks.rgb, opacity}
  const int2 *slot;         // [n_shapes] by original index: {type, slot in its type array}
  const float4 *lights;     // [4*n_lights] {pos, type} {color, cos_theta} {radial a0 a1 a2, angular a0} {dir, cull flag}
};

// Hit-point queue (one wavefront level): 48 B per entry, structure of float4 arrays, plus one
// index list per work class.  The producer files every entry under the class of its estimated
// number of live rays (reflection ray + lights in front of the surface); the consumer walks the
// lists class by class, so the 32 hit points a warp shades together need about the same number
// of ray slots.  The class is only a scheduling hint: the consumer sizes its scan from the
// exact count it computes itself.
#define RT_NCLS 4
struct HitQueue {
  float4 *r0;   // P.xyz, pixel (tile-local index, int bits)
  float4 *r1;   // N.xyz, object (original index, int bits)
  float4 *r2;   // V.xyz (direction of the ray that made the hit), throughput
  int *cls;     // [RT_NCLS][cap] entry numbers per class
  int *ctl;     // [0] entries, [1..RT_NCLS] entries per class, [5] work cursor of the consuming launch
  int cap;
};
#define RT_CTL_STRIDE 8

// Device counters (unsigned long long each).
enum {
  CT_RAYS_REFLECTION = 0, CT_RAYS_SHADOW, CT_HIT_POINTS, CT_PRIMARY_HITS,
  CT_SPH_O, CT_SPH_D, CT_PLN_O, CT_PLN_D, CT_QUAD_O, CT_QUAD_D,
  CT_LIGHT_SAMPLES, CT_CANDIDATES, CT_SLOTS, CT_COUNT
};
