/*
 * tools/gen_scene.c -- seeded generator of the synthetic benchmark scenes of
 * BASELINE.json configs 4 and 5 (recipe frozen in SURVEY.md section 8d), written in the
 * reference's scene grammar (C/parse.c:13-358) with every field under 100 characters
 * so the reference parser's fixed buffers (C/parse.c:15-17) can read them too.
 *
 *   gen_scene SPHERES PLANES QUADRICS LIGHTS SEED > out.scene
 *
 * C4: gen_scene 1024 8 0 16 1234      C5: gen_scene 2040 16 2040 16 4321
 * splitmix64 -> doubles in [0,1); all values printed with %.6f; shapes shuffled
 * (Fisher-Yates) so types interleave in file order; camera 2 x 2.
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static uint64_t g_state;
static uint64_t splitmix64(void) {
  uint64_t z = (g_state += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
static double u01(void) { return (double)(splitmix64() >> 11) * (1.0 / 9007199254740992.0); }
static double uni(double lo, double hi) { return lo + (hi - lo) * u01(); }

#define LINE 512

int main(int argc, char **argv) {
  if (argc != 6) {
    fprintf(stderr, "usage: gen_scene SPHERES PLANES QUADRICS LIGHTS SEED\n");
    return 2;
  }
  int ns = atoi(argv[1]), np = atoi(argv[2]), nq = atoi(argv[3]), nl = atoi(argv[4]);
  g_state = strtoull(argv[5], NULL, 10);
  int n = ns + np + nq;
  char (*lines)[LINE] = malloc((size_t)(n > 0 ? n : 1) * LINE);
  static const double refl_choices[5] = {0.0, 0.0, 0.25, 0.5, 0.75};
  int k = 0;

  for (int i = 0; i < ns; i++, k++) {
    double z = -uni(8, 60), x = uni(-0.9, 0.9) * fabs(z), y = uni(-0.9, 0.9) * fabs(z);
    double r = uni(0.2, 1.2);
    double kd0 = u01(), kd1 = u01(), kd2 = u01();
    double refl = refl_choices[splitmix64() % 5];
    snprintf(lines[k], LINE,
             "sphere, radius: %.6f, diffuse_color: [%.6f, %.6f, %.6f], specular_color: [1, 1, 1], "
             "position: [%.6f, %.6f, %.6f], reflectivity: %.6f, refractivity: 0, ior: 1.33",
             r, kd0, kd1, kd2, x, y, z, refl);
  }
  for (int i = 0; i < np; i++, k++) {
    /* floor-like: upward-ish unit normal, point far below the camera */
    double nx = uni(-0.25, 0.25), nz = uni(-0.25, 0.25), ny = 1.0;
    double l = sqrt(nx * nx + ny * ny + nz * nz);
    double py = -uni(20, 60);
    double kd0 = uni(0.1, 0.6), kd1 = uni(0.1, 0.6), kd2 = uni(0.1, 0.6);
    double refl = refl_choices[splitmix64() % 5];
    snprintf(lines[k], LINE,
             "plane, normal: [%.6f, %.6f, %.6f], diffuse_color: [%.6f, %.6f, %.6f], "
             "position: [0, %.6f, 0], reflectivity: %.6f",
             nx / l, ny / l, nz / l, kd0, kd1, kd2, py, refl);
  }
  for (int i = 0; i < nq; i++, k++) {
    /* axis-aligned ellipsoid ((x-cx)/rx)^2 + ((y-cy)/ry)^2 + ((z-cz)/rz)^2 = 1, expanded to a..j */
    double cz = -uni(8, 60), cx = uni(-0.9, 0.9) * fabs(cz), cy = uni(-0.9, 0.9) * fabs(cz);
    double rx = uni(0.2, 1.2), ry = uni(0.2, 1.2), rz = uni(0.2, 1.2);
    double a = 1.0 / (rx * rx), b = 1.0 / (ry * ry), c = 1.0 / (rz * rz);
    double g = -2.0 * cx * a, h = -2.0 * cy * b, ii = -2.0 * cz * c;
    double j = cx * cx * a + cy * cy * b + cz * cz * c - 1.0;
    double kd0 = u01(), kd1 = u01(), kd2 = u01();
    double refl = refl_choices[splitmix64() % 5];
    snprintf(lines[k], LINE,
             "quadric, diffuse_color: [%.6f, %.6f, %.6f], specular_color: [1, 1, 1], "
             "a: %.6f, b: %.6f, c: %.6f, d: 0, e: 0, f: 0, g: %.6f, h: %.6f, i: %.6f, j: %.6f, "
             "reflectivity: %.6f",
             kd0, kd1, kd2, a, b, c, g, h, ii, j, refl);
  }
  for (int i = n - 1; i > 0; i--) { /* shuffle shapes */
    int j = (int)(splitmix64() % (uint64_t)(i + 1));
    char tmp[LINE];
    memcpy(tmp, lines[i], LINE); memcpy(lines[i], lines[j], LINE); memcpy(lines[j], tmp, LINE);
  }

  printf("camera, width: 2, height: 2\n");
  for (int i = 0; i < n; i++) printf("%s\n", lines[i]);
  for (int i = 0; i < nl; i++) {
    double x = uni(-40, 40), y = uni(10, 60), z = -uni(0, 60);
    double c0 = uni(0.5, 2), c1 = uni(0.5, 2), c2 = uni(0.5, 2);
    printf("light, color: [%.6f, %.6f, %.6f], radial-a2: 0.002, radial-a1: 0.01, radial-a0: 0.5, "
           "position: [%.6f, %.6f, %.6f]\n", c0, c1, c2, x, y, z);
  }
  free(lines);
  return 0;
}
