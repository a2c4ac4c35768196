// Shared declarations of the tcgen05 path for the res* models (res_tc.cu, res_tc_conv.cu).
//
// Activation layout in HBM ("planar"): X[clip][cg][w][h][8] half, cg = channel group of 8
// (C padded to NCG*8), w = MFCC-coefficient axis, h = time axis (fastest).  One position is a
// 16-byte unit, so a UMMA K-major no-swizzle A operand (core matrix = 8 rows x 16 B) is simply
// 128 consecutive positions of two adjacent planes, and an im2col shift is a start-address
// offset (validated by scripts/umma_probe.cu).
#pragma once
#include <cuda.h>

#include "common.cuh"

namespace honk {

constexpr int TC_MAX_UNITS = 48;     // units (column segments of residue chains) per clip per layer
constexpr int TC_THREADS = 192;      // warp 0 TMA, warp 1 MMA, warps 2-5 epilogue

struct TcUnit {          // one work unit inside a clip
  short w_first;         // input column of packed column 0 (= first output column - d), may be < 0
  short n_cols;          // output columns in this unit
  short n_tiles;         // ceil(n_cols * Hp / 128)
  short pad;
};

struct TcLayerGeom {
  int H, W;              // feature-map size (time, coefficient)
  int d;                 // dilation
  int Hp;                // H + d rows per packed column (d zero rows on top, OOB-filled)
  int box_cols;          // packed columns per stage (max n_cols + 2)
  int n_units;           // units per clip
  int strided;           // 1: one strided TMA box per stage; 0: one box per (plane, column)
  int n_stages;
  int stage_bytes;       // box_cols * Hp * 32 (two planes)
  int stage_stride;      // stage_bytes + zero pad
  int smem_bytes;
  TcUnit units[TC_MAX_UNITS];
};

struct TcConvArgs {
  TcLayerGeom g;
  int n_clips;
  int flags;             // bit0: add residual, bit1: last layer (accumulate channel sums, no store)
  const __half* wpack;   // [9][NCG][N][8] half, B operand planes
  const float* bn_scale; // [N] 1/sqrt(var+eps)   (0 for padded channels)
  const float* bn_shift; // [N] -mean*scale
  __half* out;           // next layer's normalised input (planar)
  __half* res;           // raw residual stream (planar), read+written when bit0
  float* pool;           // [n_clips][N] channel sums (last layer)
};

int tc_conv_launch(int ncg, const TcConvArgs& a, const CUtensorMap& tmap, int grid, cudaStream_t s);
int tc_conv_prepare(int ncg, int max_smem);

}  // namespace honk
