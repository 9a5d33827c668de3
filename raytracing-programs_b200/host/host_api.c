/*
 * host/host_api.c -- small C entry points for bindings and tests: scene file ->
 * flat rt_scene through this host's own parser (no GPU involved).
 */
#include "raycast_host.h"

/* 0 ok, -1 cannot open, -2 out of memory. Parser errors exit(1) like the reference. */
int rt_host_load_scene(const char *path, rt_scene *out) {
  FILE *f = fopen(path, "r");
  if (!f) return -1;
  json_data_t js = {0};
  parse_json(f, &js);
  fclose(f);
  int r = rt_host_flatten(&js, out);
  free_shape_list(js.shapes_list);
  free_light_list(js.lights_list);
  return r == 0 ? 0 : -2;
}
