/*
 * host/scene_types.h -- host-side mirror of the reference's scene data model, so code
 * written against the reference (C/objects.h:4-73, C/parse.h:11-23) compiles unchanged
 * against this host: same type names, field names, enum values and function signatures.
 * The implementation behind them is new (scene_list.c, scene_parse.c).
 */
#ifndef RT_HOST_SCENE_TYPES_H
#define RT_HOST_SCENE_TYPES_H

#include <stdio.h>
#include <stdlib.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum { SPHERE, PLANE, QUADRIC } shape_type_t;   /* C/objects.h:4-8 */
typedef enum { POINT, SPOTLIGHT } light_type_t;         /* C/objects.h:10-13 */

/* C/objects.h:15-48. Geometry parameters share storage, one view per shape type. */
typedef struct shape_t {
  float diffuse_color[3];
  float specular_color[3];
  float position[3];
  float reflectivity;
  float refractivity;
  float ior;
  union {
    struct { float normal[3]; };                         /* PLANE   */
    struct { float radius; };                            /* SPHERE  */
    struct { float a, b, c, d, e, f, g, h, i, j; };      /* QUADRIC */
  };
  shape_type_t type;
  struct shape_t *next;
} shape_t;

/* C/objects.h:50-61 */
typedef struct light_t {
  float position[3];
  float color[3];
  float radial_coef[3];
  float theta;
  float cos_theta;
  float a0;
  float direction[3];
  light_type_t type;
  struct light_t *next;
} light_t;

/* C/parse.h:11-18 */
typedef struct json_data_t {
  float camera_width;
  float camera_height;
  shape_t *shapes_list;
  light_t *lights_list;
  int num_shapes;
  int num_lights;
} json_data_t;

extern const float PI; /* C/v3math.c:11, used by the theta conversion at C/parse.c:282 */

/* List builders, C/objects.h:63-73: append at the tail, return the (possibly new) head. */
shape_t *add_new_sphere(shape_t *head, float *diffuse, float *specular, float *position,
                        float radius, float reflectivity, float refractivity, float ior);
shape_t *add_new_plane(shape_t *head, float *diffuse, float *specular, float *position,
                       float *normal, float reflectivity);
shape_t *add_new_quadric(shape_t *head, float *diffuse, float *specular, float a, float b, float c,
                         float d, float e, float f, float g, float h, float i, float j,
                         float reflectivity);
light_t *add_new_point_light(light_t *head, float *color, float *position, float *radial_coef);
light_t *add_new_spot_light(light_t *head, float *color, float *position, float theta, float a0,
                            float *direction, float *radial_coef);
shape_t *free_shape_list(shape_t *head);   /* returns NULL */
light_t *free_light_list(light_t *head);   /* returns NULL */

/* C/parse.h:20-23 */
void parse_json(FILE *json, json_data_t *json_data);
void set_to_black(float *input);

#ifdef __cplusplus
}
#endif
#endif
