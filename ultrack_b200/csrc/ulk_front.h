// ulk_front.h -- tile-resident front end (ulk_front.cu): TMA + shared-memory staging.
#pragma once
#include "ulk_voxel.h"

// can the TMA path take these contours (dtype f32 / f16 / u8, 16-byte aligned base and row pitch)?
bool ulk_front_tile_supported(const void *ct, int ct_dtype, const UlkDims &dm);
// rank-space arrays + steepest edges in component space from the foreground words / rank prefix
int ulk_front_tile(ulk_stream_t st, const void *ct, int ct_dtype, const UlkDims &dm, i32 S, const u32 *words, const i32 *off,
                   UlkOW *ow, i32 *fgl, float *ctf, i32 *P, u32 *mw, u32 *pk, i32 *po, u32 *cmax);
