/*
 * host/ppm_out.c -- ppm_WriteOutP3 of the reference (C/ppm.c:168-184), same signature, same
 * bytes:  "P3\n" "<w> <h> \n" "<maxColor>\n"  then one decimal per line, R G B per pixel
 * (the space before the second newline is the reference's).  The reference calls fprintf
 * once per pixel (~0.2 s per Mpixel: 54 s for a 16384^2 frame); here the text is produced by
 * the GPU formatter behind rt_p3_write (csrc/rt_p3.cuh) and this file only fwrite()s the
 * pieces.  No CPU formatter exists: without a GPU the call fails loudly like raycast().
 *
 * Which pixels: the pixmap the caller passes, uploaded first -- correct for any caller.  A
 * caller that has not touched the pixmap since raycast() filled it (the reference's main,
 * C/raycast.c:57-59, and this repo's CLI) may say so with rt_host_p3_resident(1): the frame
 * raycast() left on GPU 0 is formatted directly and the upload is skipped.
 */
#include <stdlib.h>
#include <string.h>
#include "ppm_out.h"
#include "rt_abi.h"

static int g_resident = 0;
static rt_p3_stats g_p3_stats;

void rt_host_p3_resident(int yes) { g_resident = yes; }
const rt_p3_stats *rt_host_last_p3_stats(void) { return &g_p3_stats; }

static int file_sink(const void *data, size_t nbytes, void *user) {
  return fwrite(data, 1, nbytes, (FILE *)user) == nbytes ? 0 : 1;
}

void ppm_WriteOutP3(PPMFormat inData, FILE *outFile) {
  if (rt_device_count() == 0) {
    const char *g = getenv("RT_GPUS");
    if (rt_init(g ? atoi(g) : 1) != RT_OK) {
      fprintf(stderr, "Error: cannot initialise the B200 library for the P3 writer: %s\n", rt_last_error());
      exit(1);
    }
  }
  int r = RT_ERR_ARG;
  if (g_resident) r = rt_p3_write(NULL, inData.width, inData.height, inData.maxColor, file_sink, outFile, &g_p3_stats);
  if (r != RT_OK) r = rt_p3_write(inData.pixmap, inData.width, inData.height, inData.maxColor, file_sink, outFile, &g_p3_stats);
  if (r != RT_OK) {
    fprintf(stderr, "Error: P3 write failed: %s\n", rt_last_error());
    exit(1);
  }
}

/*
 * The binary writers (C/ppm.c:193-227).  The reference wrote them for an RGBA pixmap (they come
 * from its PPM converter: P6 drops every fourth byte through ppm_RemoveAlphaChannel, P7 writes
 * 4 bytes per pixel) and never calls them from the raytracer, whose pixmap is RGB.  Here
 * inData.depth says which one the caller holds: 4 = RGBA, byte for byte what the reference
 * emits; anything else = RGB as raycast() leaves it (P6 writes it as is, P7 adds an opaque
 * alpha = maxColor).  Both end with the reference's trailing newline after the raster.
 * Not part of the hot path: the frame is already in host memory and this is an fwrite.
 */
#define PPM_CHUNK_PIXELS (1u << 18)

void ppm_WriteOutP6(PPMFormat inData, FILE *outFile) {
  const size_t n = (size_t)inData.width * (size_t)inData.height;
  fprintf(outFile, "P6\n%d %d\n%d\n", inData.width, inData.height, inData.maxColor);
  if (inData.depth != 4) {
    fwrite(inData.pixmap, 1, n * 3u, outFile);
  } else {
    uint8_t *buf = (uint8_t *)malloc((size_t)PPM_CHUNK_PIXELS * 3u);
    if (buf == NULL) { fprintf(stderr, "Error: out of memory in the P6 writer\n"); exit(1); }
    for (size_t p0 = 0; p0 < n; p0 += PPM_CHUNK_PIXELS) {
      const size_t m = n - p0 < PPM_CHUNK_PIXELS ? n - p0 : PPM_CHUNK_PIXELS;
      const uint8_t *src = inData.pixmap + p0 * 4u;
      for (size_t i = 0; i < m; i++) {
        buf[3 * i] = src[4 * i]; buf[3 * i + 1] = src[4 * i + 1]; buf[3 * i + 2] = src[4 * i + 2];
      }
      fwrite(buf, 1, m * 3u, outFile);
    }
    free(buf);
  }
  fputc('\n', outFile);
}

void ppm_WriteOutP7(PPMFormat inData, FILE *outFile) {
  const size_t n = (size_t)inData.width * (size_t)inData.height;
  fprintf(outFile, "P7\nWIDTH %d\nHEIGHT %d\nDEPTH %d\nMAXVAL %d\nTUPLTYPE RGB_ALPHA\n",
          inData.width, inData.height, 4, inData.maxColor);
  if (inData.depth == 4) {
    fwrite(inData.pixmap, 1, n * 4u, outFile);
  } else {
    uint8_t *buf = (uint8_t *)malloc((size_t)PPM_CHUNK_PIXELS * 4u);
    if (buf == NULL) { fprintf(stderr, "Error: out of memory in the P7 writer\n"); exit(1); }
    for (size_t p0 = 0; p0 < n; p0 += PPM_CHUNK_PIXELS) {
      const size_t m = n - p0 < PPM_CHUNK_PIXELS ? n - p0 : PPM_CHUNK_PIXELS;
      const uint8_t *src = inData.pixmap + p0 * 3u;
      for (size_t i = 0; i < m; i++) {
        buf[4 * i] = src[3 * i]; buf[4 * i + 1] = src[3 * i + 1]; buf[4 * i + 2] = src[3 * i + 2];
        buf[4 * i + 3] = inData.maxColor;
      }
      fwrite(buf, 1, m * 4u, outFile);
    }
    free(buf);
  }
  fputc('\n', outFile);
}

float ppm_clamp(float value, float lower_bound, float upper_bound) {
  if (value > upper_bound) value = upper_bound;
  if (value < lower_bound) value = lower_bound;
  return value;
}
