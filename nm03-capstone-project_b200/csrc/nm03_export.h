// nm03_export.h — GPU half of the host-side JPEG export: the grey rendering of each original slice
// (ImageRenderer's window/level on a canvas of the slice's own size), level-shifted, transformed and quantised per 8x8
// block for a baseline JPEG.  The host keeps the entropy coding (host/jpeg.hpp encode_blocks).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace nm03 {

struct ExportParams {
  const void* in;             // [n][h][w] int16 / uint16, the slices the pipeline ran on
  int n, w, h;                // w, h multiples of 8
  int input_unsigned;
  int2* mm;                   // scratch [n]: smallest and largest sample of each slice (the renderer's window)
  int16_t* blocks;            // [n][(h/8)*(w/8)][64] quantised coefficients, zigzag order, blocks in raster order
  unsigned long long* nz;     // [n][(h/8)*(w/8)] bit k set iff coefficient k of the block is non-zero
  float rq[64];               // reciprocal quantiser incl. the transform's output scaling, natural order (jpeg::make_tables)
};

cudaError_t launch_export_blocks(const ExportParams& p, cudaStream_t s);

// The same for the rendering of the RESULT: SegmentationRenderer of a {0,1} label image on a canvas of its own size
// (host/nm03.hpp paintBits): labelled pixels in the interior grey, labelled pixels within `radius` (disc; outside the
// image counts as unlabelled) of an unlabelled one in the border grey, everything else the background grey.
struct ExportSegParams {
  const uint32_t* bits;       // [n][h][wpr] the result mask, bit x%32 of word x/32 = pixel x (nm03_outputs.dilated_bits)
  int n, w, h, wpr;           // w, h multiples of 8
  int radius;                 // border radius, 0..7
  int8_t hw[15];              // half-width of the border disc at row offset b (index b + radius)
  uint8_t background, interior, border;
  int16_t* blocks;
  unsigned long long* nz;
  float rq[64];
};
cudaError_t launch_export_seg_blocks(const ExportSegParams& p, cudaStream_t s);

}  // namespace nm03
