/*
 * host/raycast_host.h -- the reference's operator for the hot path, same signature
 * (C/raycast.h:8).  Behind it sits the B200 library of include/rt_abi.h.
 */
#ifndef RT_HOST_RAYCAST_H
#define RT_HOST_RAYCAST_H
#include "ppm_out.h"
#include "scene_types.h"
#include "rt_abi.h"
#ifdef __cplusplus
extern "C" {
#endif

/* Render json_struct into photo_data.pixmap (RGB8, row-major, row 0 = top).
 * Like the reference (C/raycast.c:104-107) it consumes both lists: they are freed and
 * the heads set to NULL (the reference leaves them dangling at its own stack arrays).
 * Any library failure prints "Error: ..." to stderr and exit(1)s, the reference's
 * only error convention. */
void raycast(json_data_t *json_struct, PPMFormat photo_data);

/* Linked lists -> the flat file-order arrays of rt_abi.h (the copy of
 * C/raycast.c:87-102). Returns 0 or -1 (out of memory). Caller frees with
 * rt_host_free_scene. Lists are left untouched. */
int rt_host_flatten(const json_data_t *json_struct, rt_scene *out);
void rt_host_free_scene(rt_scene *scene);

/* Stats of the last raycast() call (for the optional RT_VERBOSE line of main). */
const rt_stats *rt_host_last_stats(void);

#ifdef __cplusplus
}
#endif
#endif
