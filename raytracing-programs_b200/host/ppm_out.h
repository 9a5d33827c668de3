/*
 * host/ppm_out.h -- the slice of the reference's PPM module that the raytracer uses
 * (C/ppm.h:6-12,17,25): the PPMFormat descriptor, the ASCII P3 writer and ppm_clamp.
 * Same names and signatures so main()/raycast() read like the reference's.
 */
#ifndef RT_HOST_PPM_OUT_H
#define RT_HOST_PPM_OUT_H
#include <stdint.h>
#include <stdio.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct PPMFormat {          /* C/ppm.h:6-12 */
  int width, height, size;          /* size = 3*width*height; overflows int above ~26.7k^2, */
  uint8_t maxColor;                 /* so this host never trusts it and recomputes in size_t */
  uint8_t depth;
  char *tupleType;
  uint8_t *pixmap;
} PPMFormat;

void ppm_WriteOutP3(PPMFormat inData, FILE *outFile);   /* C/ppm.c:168-184 */
/* The pixmap passed to the next ppm_WriteOutP3 calls is exactly what raycast() rendered last
 * (untouched since): format the frame still resident on GPU 0 instead of uploading it. */
void rt_host_p3_resident(int yes);
/* Binary writers; inData.depth == 4: RGBA pixmap (the reference's assumption), else RGB. */
void ppm_WriteOutP6(PPMFormat inData, FILE *outFile);   /* C/ppm.c:193-203 */
void ppm_WriteOutP7(PPMFormat inData, FILE *outFile);   /* C/ppm.c:212-227 */
float ppm_clamp(float value, float lower_bound, float upper_bound); /* C/ppm.c:350-359 */

#ifdef __cplusplus
}
#endif
#endif
