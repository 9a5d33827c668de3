/*
 * tools/scene_dump.c -- parse a .scene with this host's parser and print what came out
 * (counts, then every float as hex bits). No GPU needed; used by the parser tests,
 * including the error paths (the parser exit(1)s with the reference's messages).
 */
#include <stdint.h>
#include <string.h>
#include "scene_types.h"

static void hex(const char *tag, const float *v, int n) {
  printf("%s", tag);
  for (int i = 0; i < n; i++) { uint32_t u; memcpy(&u, &v[i], 4); printf(" %08x", u); }
  printf("\n");
}

int main(int argc, char **argv) {
  if (argc != 2) { fprintf(stderr, "usage: scene_dump FILE.scene\n"); return 2; }
  FILE *f = fopen(argv[1], "r");
  if (!f) { fprintf(stderr, "Error: Unable to open the input scene file: %s\n", argv[1]); return 1; }
  json_data_t js = {0};
  parse_json(f, &js);
  fclose(f);
  printf("camera %a %a shapes %d lights %d\n", js.camera_width, js.camera_height, js.num_shapes, js.num_lights);
  for (shape_t *s = js.shapes_list; s; s = s->next) {
    printf("shape type %d\n", (int)s->type);
    hex(" kd", s->diffuse_color, 3); hex(" ks", s->specular_color, 3);
    if (s->type != QUADRIC) hex(" pos", s->position, 3);
    hex(" rri", &s->reflectivity, 3);
    hex(" geo", &s->a, s->type == SPHERE ? 1 : s->type == PLANE ? 3 : 10);
  }
  for (light_t *l = js.lights_list; l; l = l->next) {
    printf("light type %d\n", (int)l->type);
    hex(" pos", l->position, 3); hex(" col", l->color, 3); hex(" rad", l->radial_coef, 3);
    if (l->type == SPOTLIGHT) { hex(" tca", &l->theta, 3); hex(" dir", l->direction, 3); }
  }
  return 0;
}
