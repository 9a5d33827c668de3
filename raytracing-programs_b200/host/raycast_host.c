/*
 * host/raycast_host.c -- raycast(json_data_t*, PPMFormat) of the reference
 * (C/raycast.h:8; body C/raycast.c:79-130) re-hosted on the B200 library: flatten the
 * lists in file order, upload, render on the GPU(s), pixels land in photo_data.pixmap.
 * No CPU rendering path exists here: without a GPU the call fails loudly.
 * Environment knobs (the CLI stays 4 positional arguments): RT_GPUS (default 1,
 * 0 = all visible), RT_BAND_ROWS (band height of the multi-GPU row interleave), RT_MAX_DEPTH, RT_MODE=shoot.
 */
#include <string.h>
#include "raycast_host.h"

static rt_stats g_last_stats;

int rt_host_flatten(const json_data_t *js, rt_scene *out) {
  memset(out, 0, sizeof *out);
  rt_shape *shapes = (rt_shape *)calloc((size_t)js->num_shapes + 1, sizeof(rt_shape));
  rt_light *lights = (rt_light *)calloc((size_t)js->num_lights + 1, sizeof(rt_light));
  if (!shapes || !lights) { free(shapes); free(lights); return -1; }
  int n = 0;
  for (const shape_t *s = js->shapes_list; s && n < js->num_shapes; s = s->next, n++) {
    rt_shape *d = &shapes[n];
    d->type = (int32_t)s->type;
    memcpy(d->diffuse, s->diffuse_color, sizeof d->diffuse);
    memcpy(d->specular, s->specular_color, sizeof d->specular);
    d->reflectivity = s->reflectivity;
    d->refractivity = s->refractivity;
    d->ior = s->ior;
    if (s->type == SPHERE) {
      memcpy(d->position, s->position, sizeof d->position);
      d->params[0] = s->radius;
    } else if (s->type == PLANE) {
      memcpy(d->position, s->position, sizeof d->position);
      memcpy(d->params, s->normal, 3 * sizeof(float));
    } else {
      memcpy(d->params, &s->a, 10 * sizeof(float));
    }
  }
  out->num_shapes = n;
  n = 0;
  for (const light_t *l = js->lights_list; l && n < js->num_lights; l = l->next, n++) {
    rt_light *d = &lights[n];
    d->type = (int32_t)l->type;
    memcpy(d->position, l->position, sizeof d->position);
    memcpy(d->color, l->color, sizeof d->color);
    memcpy(d->radial, l->radial_coef, sizeof d->radial);
    if (l->type == SPOTLIGHT) {
      d->theta = l->theta;
      d->cos_theta = l->cos_theta;
      d->a0 = l->a0;
      memcpy(d->direction, l->direction, sizeof d->direction);
    }
  }
  out->num_lights = n;
  out->camera_width = js->camera_width;
  out->camera_height = js->camera_height;
  out->shapes = shapes;
  out->lights = lights;
  return 0;
}

void rt_host_free_scene(rt_scene *scene) {
  free((void *)scene->shapes);
  free((void *)scene->lights);
  scene->shapes = NULL;
  scene->lights = NULL;
}

const rt_stats *rt_host_last_stats(void) { return &g_last_stats; }

static void fail(const char *what) {
  fprintf(stderr, "Error: %s: %s\n", what, rt_last_error());
  exit(1);
}

void raycast(json_data_t *json_struct, PPMFormat photo_data) {
  rt_scene scene;
  if (rt_host_flatten(json_struct, &scene) != 0) {
    fprintf(stderr, "Error: out of memory while flattening the scene.\n");
    exit(1);
  }
  json_struct->lights_list = free_light_list(json_struct->lights_list);
  json_struct->shapes_list = free_shape_list(json_struct->shapes_list);

  if (rt_device_count() == 0) {
    const char *g = getenv("RT_GPUS");
    if (rt_init(g ? atoi(g) : 1) != RT_OK) fail("cannot initialise the B200 render library");
  }
  const char *md = getenv("RT_MAX_DEPTH");
  if (md && rt_set_max_depth(atoi(md)) != RT_OK) fail("bad RT_MAX_DEPTH");
  const char *mode = getenv("RT_MODE");   /* "shoot": the reference's dormant recursive tracer, C/raycast.c:152-296 */
  if (mode && rt_set_mode(strcmp(mode, "shoot") == 0 ? RT_MODE_SHOOT : RT_MODE_ITERATIVE) != RT_OK) fail("bad RT_MODE");
  if (rt_scene_upload(&scene) != RT_OK) fail("scene upload failed");
  if (rt_render(photo_data.width, photo_data.height, photo_data.pixmap, &g_last_stats) != RT_OK)
    fail("render failed");
  rt_host_free_scene(&scene);
}
