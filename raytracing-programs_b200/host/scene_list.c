/*
 * host/scene_list.c -- list builders behind the reference's add_new_* / free_* API
 * (C/objects.h:63-73; behaviour of C/objects.c:20-221): tail append in file order,
 * plane/quadric refractivity 0 and ior 1, cos_theta = cos(theta) in double.
 * New implementation: iterative (the reference recurses once per node, which is a
 * stack hazard on the 4096-object synthetic scenes), zero-initialised nodes, and the tail of
 * the list being built is remembered, so appending n objects is O(n) instead of the
 * reference's O(n^2) walk (C/objects.c:40).
 */
#include <math.h>
#include <string.h>
#include "scene_types.h"

const float PI = 3.141592654f; /* C/v3math.c:11 */

static shape_t *shape_node(shape_type_t type, const float *diffuse, const float *specular,
                           float reflectivity) {
  shape_t *n = (shape_t *)calloc(1, sizeof(shape_t));
  if (!n) { fprintf(stderr, "Error: out of memory while building the shape list.\n"); exit(1); }
  memcpy(n->diffuse_color, diffuse, sizeof n->diffuse_color);
  memcpy(n->specular_color, specular, sizeof n->specular_color);
  n->reflectivity = reflectivity;
  n->refractivity = 0;
  n->ior = 1;
  n->type = type;
  return n;
}

/* Tail cache: valid while the same head keeps being appended to; dropped by free_*_list. */
static shape_t *g_shape_head, *g_shape_tail;
static light_t *g_light_head, *g_light_tail;

static shape_t *shape_append(shape_t *head, shape_t *n) {
  if (!head) { g_shape_head = g_shape_tail = n; return n; }
  shape_t *t = (head == g_shape_head && g_shape_tail && !g_shape_tail->next) ? g_shape_tail : head;
  while (t->next) t = t->next;
  t->next = n;
  g_shape_head = head; g_shape_tail = n;
  return head;
}

shape_t *add_new_sphere(shape_t *head, float *diffuse, float *specular, float *position,
                        float radius, float reflectivity, float refractivity, float ior) {
  shape_t *n = shape_node(SPHERE, diffuse, specular, reflectivity);
  memcpy(n->position, position, sizeof n->position);
  n->refractivity = refractivity;
  n->ior = ior;
  n->radius = radius;
  return shape_append(head, n);
}

shape_t *add_new_plane(shape_t *head, float *diffuse, float *specular, float *position,
                       float *normal, float reflectivity) {
  shape_t *n = shape_node(PLANE, diffuse, specular, reflectivity);
  memcpy(n->position, position, sizeof n->position);
  memcpy(n->normal, normal, sizeof n->normal);
  return shape_append(head, n);
}

shape_t *add_new_quadric(shape_t *head, float *diffuse, float *specular, float a, float b, float c,
                         float d, float e, float f, float g, float h, float i, float j,
                         float reflectivity) {
  shape_t *n = shape_node(QUADRIC, diffuse, specular, reflectivity);
  const float coef[10] = {a, b, c, d, e, f, g, h, i, j};
  memcpy(&n->a, coef, sizeof coef);
  return shape_append(head, n);
}

static light_t *light_append(light_t *head, light_t *n) {
  if (!head) { g_light_head = g_light_tail = n; return n; }
  light_t *t = (head == g_light_head && g_light_tail && !g_light_tail->next) ? g_light_tail : head;
  while (t->next) t = t->next;
  t->next = n;
  g_light_head = head; g_light_tail = n;
  return head;
}

static light_t *light_node(light_type_t type, const float *color, const float *position,
                           const float *radial_coef) {
  light_t *n = (light_t *)calloc(1, sizeof(light_t));
  if (!n) { fprintf(stderr, "Error: out of memory while building the light list.\n"); exit(1); }
  memcpy(n->color, color, sizeof n->color);
  memcpy(n->position, position, sizeof n->position);
  memcpy(n->radial_coef, radial_coef, sizeof n->radial_coef);
  n->type = type;
  return n;
}

light_t *add_new_point_light(light_t *head, float *color, float *position, float *radial_coef) {
  return light_append(head, light_node(POINT, color, position, radial_coef));
}

light_t *add_new_spot_light(light_t *head, float *color, float *position, float theta, float a0,
                            float *direction, float *radial_coef) {
  light_t *n = light_node(SPOTLIGHT, color, position, radial_coef);
  memcpy(n->direction, direction, sizeof n->direction);
  n->a0 = a0;
  n->theta = theta;
  n->cos_theta = (float)cos((double)theta); /* C/objects.c:179 */
  return light_append(head, n);
}

shape_t *free_shape_list(shape_t *head) {
  g_shape_head = g_shape_tail = NULL;
  while (head) { shape_t *nx = head->next; free(head); head = nx; }
  return NULL;
}

light_t *free_light_list(light_t *head) {
  g_light_head = g_light_tail = NULL;
  while (head) { light_t *nx = head->next; free(head); head = nx; }
  return NULL;
}

void set_to_black(float *input) { input[0] = input[1] = input[2] = 0; }
