#define ORBX_MAX_DYN_SMEM (227 * 1024)   /* sm_100a: opt-in dynamic shared memory per CTA */
// orbx_internal.h -- structures shared by the host runtime (orbx_api.cu) and the kernels.
//
// Data layout in HBM (per pipeline slot, B = max_batch frames; all sizes for one image shape):
//   img0     u8   [B][stride0]               level 0 = the input frames (device staging of the
//                                            H2D copy, or the caller's device buffer), pitch0
//   pyr      u8   [B][frame_stride]          levels 1.. packed per frame; level l at lvl[l].off,
//                                            row pitch lvl[l].pitch; NO border is materialised
//                                            (nothing on the path reads it)
//   cellcnt  u32  [B][cells_total]           post-NMS corner count per 30-px FAST cell
//   cellkeys u32  [B][keys_total]            per cell a fixed slot of lvl[l].cell_cap keys
//   keysA/B  u32  [B][keys_total]            quadtree ping-pong buffers (spill when a level's
//                                            keys do not fit shared memory)
//   surv     u32  [B][surv_total]            quadtree survivors per level (list order)
//   survcnt  u32  [B][ORBX_MAX_LEVELS]
//   dest     u32  [B][surv_total]            final output row of each survivor (lapping layout)
//   header   i32  [B][4]                     {n, mono_index, status, 0}
//   kps      orbx_keypoint [B][out_cap]
//   desc     u8   [B][out_cap][32]
// A key packs one FAST corner: x | y << 12 | response << 24, (x, y) relative to (16,16).
#ifndef ORBX_INTERNAL_H
#define ORBX_INTERNAL_H

#include <stdint.h>
#include <vector_types.h>

#include "../../include/orbx.h"

#define ORBX_EDGE 19        // EDGE_THRESHOLD, ORBextractor.cc:71
#define ORBX_FAST_BORDER 16 // EDGE_THRESHOLD - 3, ORBextractor.cc:771
#define ORBX_KEY_X(k) ((int)((k) & 0xfffu))
#define ORBX_KEY_Y(k) ((int)(((k) >> 12) & 0xfffu))
#define ORBX_KEY_R(k) ((int)((k) >> 24))
#define ORBX_MAKE_KEY(x, y, r) ((uint32_t)(x) | ((uint32_t)(y) << 12) | ((uint32_t)(r) << 24))
#define ORBX_MAX_DIM 4096   // keys hold 12-bit coordinates

struct OrbxLevel {
    int w, h;            // level size (cvRound(cols * mvInvScaleFactor[l]), ORBextractor.cc:1156-1157)
    int pitch;           // bytes per row in the slab
    int _pad0;
    long long off;       // byte offset of the level inside a frame's slab
    // FAST cell grid (ORBextractor.cc:771-785)
    int n_cols, n_rows;  // cells
    uint32_t n_cols_inv; // ceil(2^32 / n_cols), 0 for n_cols == 1: c / n_cols == umulhi(c, inv) for c < 2^16
    int _pad2;
    int w_cell, h_cell;
    int max_bx, max_by;  // maxBorderX/Y = w - 16, h - 16
    int cell_cap;        // ceil(w_cell/2) * ceil(h_cell/2): NMS survivors cannot be 8-adjacent
    int cell_off;        // first cell of this level in cellcnt[]
    long long key_off;   // first key slot of this level in cellkeys[] / keysA / keysB
    int key_cap;         // n_cols * n_rows * cell_cap
    // quadtree (ORBextractor.cc:537-568)
    int n_target;        // mnFeaturesPerLevel[l]
    int n_ini;           // round(W / H)
    float hx;            // float(W) / nIni
    int node_cap;        // upper bound on list nodes
    int surv_off;        // first survivor slot of this level in surv[] / dest[]
    int surv_cap;        // node_cap
    // resize tables for producing this level from level l-1 (device pointers)
    const uint16_t* xofs;  // [w]  source column of the first tap
    const short* xa;       // [2*w] 11-bit coefficients (a0, a1) interleaved
    const uint16_t* yofs;  // [h]
    const short* ya;       // [2*h]
    const uint2* ytab;     // [h]  {yofs, b0 | b1 << 16}: one 8-byte load per output row
    int x_step_max;        // max over x of xofs[x+1] - xofs[x] (the fast kernel needs <= 3)
    int y_monotone;        // 1 if ytab's source rows strictly increase (the fast kernel needs it)
    float scale;         // mvScaleFactor[l]
    float size;          // (int)(31 * mvScaleFactor[l]) as float
};

struct OrbxPlan {
    int nlevels;
    int w, h;
    int ini_th, min_th;
    int cells_total;
    int surv_total;
    int out_cap;
    long long keys_total;
    long long frame_stride;  // bytes of one frame's pyramid slab
    int max_tile_w, max_tile_h;  // FAST tile (cell + 6) maxima over levels
    int _pad;
    // one record per FAST cell (all levels, cell_off order), built once per shape on the host:
    //   x = iniX | iniY << 16          tile origin in the level (ORBextractor.cc:787-804)
    //   y = tile_w | tile_h << 8 | level << 16     (tile_w == 0: the reference skips this cell)
    //   z = first key slot of the cell inside a frame's cellkeys[]
    //   w = xadd | yadd << 16          cell origin relative to the detection origin (16,16), + 3
    const uint4* cell_tab;
    OrbxLevel lvl[ORBX_MAX_LEVELS];
};

// Level 0 is the caller's image itself (never copied into the slab): device pointer to the
// batch, row pitch and frame stride in bytes.  Pointer and pitch are 4-byte aligned.
struct OrbxSrc {
    const uint8_t* img0;
    long long stride0;
    int pitch0;
    int _pad;
};

// Result sink of the fused multi-GPU gather (orbx_set_result_sink): k_describe writes every result
// row of frame f also to frame first_frame + f of these arrays (same capacity per frame as the local
// slab), which live on the gathering GPU -- local memory, or peer memory mapped over NVLink.
struct OrbxSinkArgs {
    int32_t* hdr;          // [frames][4], nullptr = no sink
    orbx_keypoint* kps;    // [frames][out_cap]
    uint8_t* desc;         // [frames][out_cap][32]
    int first_frame;
    int _pad;
    // completion signal of the fused gather: after the batch's rows are in the sink, the producer bumps its own
    // device-side batch counter and publishes it at `flag` (a word of the sink, system scope); with back-pressure it
    // first waits until the consumer has released the previous batch at `ack` (nullptr: no back-pressure)
    uint32_t* flag;
    const uint32_t* ack;
    uint32_t* epoch_dev;   // this slot's batch counter (local device memory)
    uint32_t* err_dev;     // set to 1 when the back-pressure wait timed out
};

#endif  // ORBX_INTERNAL_H
