/*
 * host/scene_parse.c -- clean-room reader for the reference's ".scene" grammar, behind
 * the reference's own entry point parse_json(FILE*, json_data_t*) (C/parse.h:20,
 * behaviour of C/parse.c:13-358).  One object per line:
 *     TYPE, key: value, key: [a, b, c], ...
 * TYPE is camera | sphere | plane | quadric | light.  Quirks kept on purpose, because
 * they decide which files are accepted and which floats come out:
 *   - spaces are dropped from keys, leading spaces from values (C/parse.c:402-430);
 *   - unknown keys are ignored and not counted (e.g. "radius" on a plane);
 *   - a recognised key counts every time it appears; totals must be exactly
 *     camera 2, sphere 7, plane 5, quadric 13, light 5 (point) or 8 (spot)
 *     (C/parse.c:59,108,162,244,307); a missing diffuse/specular colour is black and
 *     counts as present (:99-106);
 *   - "theta: 0" makes a point light and does not count (:277-279); any other theta is
 *     scaled by 180.0/PI in double with the float PI (:282) -- the reference's inverted
 *     degree/radian conversion is part of the image;
 *   - values are converted with sscanf("%f") / sscanf("[%f, %f, %f]") exactly as the
 *     reference does, and scratch values persist from line to line;
 *   - reflectivity/refractivity outside [0,1] and a camera count != 1 are fatal;
 *     all failures print the reference's message to stderr and exit(1).
 * Unlike the reference there is no 100-byte field limit, no recursion and no read past
 * the end of a field that lacks ':' (such a field is ignored).
 * Table-driven: each object type is a list of {key, kind, destination}.
 */
#include <string.h>
#include "scene_types.h"

enum { K_SCALAR, K_VEC3, K_THETA };

typedef struct field_spec {
  const char *key;
  int kind;
  float *dst;
  int *seen; /* optional presence flag */
} field_spec;

/* Scratch that deliberately survives from one line to the next (see header). */
typedef struct scratch {
  float color[3], diffuse[3], specular[3], pos[3], norm[3], direction[3], radial[3];
  float reflectivity, refractivity, ior, radius, theta, a0;
  float quad[10];
} scratch;

static void die(const char *msg) {
  fprintf(stderr, "%s", msg);
  exit(1);
}

/* Next top-level field of [p, end): stops at a comma outside [...]; returns the start
 * of the following field (or end). The field is NUL-terminated in place. */
static char *next_field(char *p, char *end, char **field) {
  *field = p;
  int depth = 0;
  while (p < end) {
    if (*p == '[') depth = 1;
    else if (*p == ']') depth = 0;
    else if (*p == ',' && !depth) { *p = '\0'; return p + 1; }
    p++;
  }
  return end;
}

/* "  key name : value" -> key with every space removed, value without leading spaces. */
static int split_field(char *field, char **key, char **value) {
  char *colon = strchr(field, ':');
  if (!colon) return 0;
  char *w = field;
  for (char *r = field; r < colon; r++)
    if (*r != ' ') *w++ = *r;
  *w = '\0';
  char *v = colon + 1;
  while (*v == ' ') v++;
  *key = field;
  *value = v;
  return 1;
}

/* Apply every field of the line to the table; returns the number of counted fields. */
static int apply_fields(char *p, char *end, const field_spec *tab, int ntab, scratch *sc,
                        int *is_spot) {
  int counted = 0;
  while (p < end) {
    char *field, *key, *value;
    p = next_field(p, end, &field);
    if (!split_field(field, &key, &value)) continue;
    for (int i = 0; i < ntab; i++) {
      if (strcmp(key, tab[i].key) != 0) continue;
      if (tab[i].kind == K_VEC3) {
        sscanf(value, "[%f, %f, %f]", &tab[i].dst[0], &tab[i].dst[1], &tab[i].dst[2]);
      } else {
        sscanf(value, "%f", tab[i].dst);
        if (tab[i].kind == K_THETA) {
          if (sc->theta == 0.0) break;                       /* point light, not counted */
          sc->theta = sc->theta * (180.0 / PI);              /* double product, float store */
          *is_spot = 1;
        }
      }
      if (tab[i].seen) *tab[i].seen = 1;
      counted++;
      break;
    }
  }
  return counted;
}

#define NTAB(t) ((int)(sizeof(t) / sizeof((t)[0])))

void parse_json(FILE *json, json_data_t *jd) {
  /* slurp the stream */
  size_t cap = 1 << 16, len = 0;
  char *buf = (char *)malloc(cap);
  if (!buf) die("Error: out of memory while reading the scene.\n");
  for (;;) {
    if (cap - len < 2) {
      cap *= 2;
      buf = (char *)realloc(buf, cap);
      if (!buf) die("Error: out of memory while reading the scene.\n");
    }
    size_t got = fread(buf + len, 1, cap - len - 1, json);
    len += got;
    if (got == 0) break;
  }
  buf[len] = '\0';

  scratch sc;
  memset(&sc, 0, sizeof sc);
  int cameras = 0;
  char *p = buf, *file_end = buf + len;

  while (p < file_end) {
    char *eol = memchr(p, '\n', (size_t)(file_end - p));
    char *end = eol ? eol : file_end;
    *end = '\0';
    char *type;
    char *rest = next_field(p, end, &type);
    int seen_diffuse = 0, seen_specular = 0, is_spot = 0, n;

    if (strcmp(type, "camera") == 0) {
      const field_spec tab[] = {{"width", K_SCALAR, &jd->camera_width, 0},
                                {"height", K_SCALAR, &jd->camera_height, 0}};
      cameras++;
      n = apply_fields(rest, end, tab, NTAB(tab), &sc, &is_spot);
      if (n != 2) die("Error: A camera width or height was not given.\n");
    } else if (strcmp(type, "sphere") == 0) {
      const field_spec tab[] = {{"diffuse_color", K_VEC3, sc.diffuse, &seen_diffuse},
                                {"specular_color", K_VEC3, sc.specular, &seen_specular},
                                {"position", K_VEC3, sc.pos, 0},
                                {"radius", K_SCALAR, &sc.radius, 0},
                                {"reflectivity", K_SCALAR, &sc.reflectivity, 0},
                                {"refractivity", K_SCALAR, &sc.refractivity, 0},
                                {"ior", K_SCALAR, &sc.ior, 0}};
      n = apply_fields(rest, end, tab, NTAB(tab), &sc, &is_spot);
      if (!seen_specular) { set_to_black(sc.specular); n++; }
      if (!seen_diffuse) { set_to_black(sc.diffuse); n++; }
      if (n != 7) die("Error: A field for a sphere was missing.\n");
      if (sc.reflectivity < 0 || sc.reflectivity > 1)
        die("Error: invalid reflectivity for a sphere. Must be between 0 and 1.\n");
      if (sc.refractivity < 0 || sc.refractivity > 1)
        die("Error: invalid refractivity for a sphere. Must be between 0 and 1.\n");
      jd->shapes_list = add_new_sphere(jd->shapes_list, sc.diffuse, sc.specular, sc.pos, sc.radius,
                                       sc.reflectivity, sc.refractivity, sc.ior);
      jd->num_shapes += 1;
    } else if (strcmp(type, "plane") == 0) {
      const field_spec tab[] = {{"diffuse_color", K_VEC3, sc.diffuse, &seen_diffuse},
                                {"specular_color", K_VEC3, sc.specular, &seen_specular},
                                {"position", K_VEC3, sc.pos, 0},
                                {"normal", K_VEC3, sc.norm, 0},
                                {"reflectivity", K_SCALAR, &sc.reflectivity, 0}};
      n = apply_fields(rest, end, tab, NTAB(tab), &sc, &is_spot);
      if (!seen_specular) { set_to_black(sc.specular); n++; }
      if (!seen_diffuse) { set_to_black(sc.diffuse); n++; }
      if (n != 5) die("Error: A field for a plane was missing.\n");
      if (sc.reflectivity < 0 || sc.reflectivity > 1) /* the reference says "sphere" here too */
        die("Error: invalid reflectivity for a sphere. Must be between 0 and 1.\n");
      jd->shapes_list = add_new_plane(jd->shapes_list, sc.diffuse, sc.specular, sc.pos, sc.norm,
                                      sc.reflectivity);
      jd->num_shapes += 1;
    } else if (strcmp(type, "quadric") == 0) {
      const field_spec tab[] = {{"diffuse_color", K_VEC3, sc.diffuse, &seen_diffuse},
                                {"specular_color", K_VEC3, sc.specular, &seen_specular},
                                {"reflectivity", K_SCALAR, &sc.reflectivity, 0},
                                {"a", K_SCALAR, &sc.quad[0], 0}, {"b", K_SCALAR, &sc.quad[1], 0},
                                {"c", K_SCALAR, &sc.quad[2], 0}, {"d", K_SCALAR, &sc.quad[3], 0},
                                {"e", K_SCALAR, &sc.quad[4], 0}, {"f", K_SCALAR, &sc.quad[5], 0},
                                {"g", K_SCALAR, &sc.quad[6], 0}, {"h", K_SCALAR, &sc.quad[7], 0},
                                {"i", K_SCALAR, &sc.quad[8], 0}, {"j", K_SCALAR, &sc.quad[9], 0}};
      n = apply_fields(rest, end, tab, NTAB(tab), &sc, &is_spot);
      if (!seen_specular) { set_to_black(sc.specular); n++; }
      if (!seen_diffuse) { set_to_black(sc.diffuse); n++; }
      if (n != 13) die("Error: A field for a quadric was missing.\n");
      if (sc.reflectivity < 0 || sc.reflectivity > 1)
        die("Error: invalid reflectivity for a sphere. Must be between 0 and 1.\n");
      jd->shapes_list = add_new_quadric(jd->shapes_list, sc.diffuse, sc.specular, sc.quad[0],
                                        sc.quad[1], sc.quad[2], sc.quad[3], sc.quad[4], sc.quad[5],
                                        sc.quad[6], sc.quad[7], sc.quad[8], sc.quad[9],
                                        sc.reflectivity);
      jd->num_shapes += 1;
    } else if (strcmp(type, "light") == 0) {
      const field_spec tab[] = {{"color", K_VEC3, sc.color, 0},
                                {"position", K_VEC3, sc.pos, 0},
                                {"theta", K_THETA, &sc.theta, 0},
                                {"radial-a0", K_SCALAR, &sc.radial[0], 0},
                                {"radial-a1", K_SCALAR, &sc.radial[1], 0},
                                {"radial-a2", K_SCALAR, &sc.radial[2], 0},
                                {"angular-a0", K_SCALAR, &sc.a0, 0},
                                {"direction", K_VEC3, sc.direction, 0}};
      n = apply_fields(rest, end, tab, NTAB(tab), &sc, &is_spot);
      if ((!is_spot && n != 5) || (is_spot && n != 8)) die("Error: A field for a light was missing.\n");
      if (is_spot)
        jd->lights_list = add_new_spot_light(jd->lights_list, sc.color, sc.pos, sc.theta, sc.a0,
                                             sc.direction, sc.radial);
      else
        jd->lights_list = add_new_point_light(jd->lights_list, sc.color, sc.pos, sc.radial);
      jd->num_lights += 1;
    }
    /* any other first token (blank line, comment, unknown type): ignored */
    p = end + 1;
  }
  free(buf);
  if (cameras != 1) die("Error: The scene must have one and only one camera.\n");
}
