// host_capi.cpp — C entry points over the host-side helpers (DICOM reader/writer, phantom, JPEG,
// renderer restatement) so that they can be exercised from tests without a GPU (libnm03_host.so).
#include <cstring>

#include "dicom.hpp"
#include "jpeg.hpp"
#include "phantom.hpp"
#include "nm03.hpp"

extern "C" {

// returns 0 or -1 (message in err)
int nm03h_dicom_write(const char* path, const uint16_t* px, int cols, int rows, int is_signed, int instance, int explicit_vr,
                      char* err, int errlen) {
  try { nm03::dicom::save(path, px, cols, rows, is_signed != 0, instance, explicit_vr != 0); return 0; }
  catch (const std::exception& e) { if (err) std::strncpy(err, e.what(), errlen - 1); return -1; }
}

// same, with RescaleSlope / RescaleIntercept written as the given decimal strings
int nm03h_dicom_write_rescale(const char* path, const uint16_t* px, int cols, int rows, int is_signed, int instance, int explicit_vr,
                              const char* slope, const char* intercept, char* err, int errlen) {
  try { nm03::dicom::save(path, px, cols, rows, is_signed != 0, instance, explicit_vr != 0, slope, intercept); return 0; }
  catch (const std::exception& e) { if (err) { std::strncpy(err, e.what(), errlen - 1); err[errlen - 1] = 0; } return -1; }
}

// header through the positional reader the batch driver uses (first 16 KiB, the whole file only when needed)
int nm03h_dicom_view(const char* path, int* cols, int* rows, long* pixel_offset, char* err, int errlen) {
  try {
    nm03::dicom::FileReader r(path);
    const nm03::dicom::View v = r.view();
    *cols = v.cols; *rows = v.rows; *pixel_offset = (long)v.pixel_offset;
    return 0;
  } catch (const std::exception& e) { if (err) { std::strncpy(err, e.what(), errlen - 1); err[errlen - 1] = 0; } return -1; }
}

// first call with px = NULL to get the size
int nm03h_dicom_read(const char* path, int* cols, int* rows, int* is_signed, int* instance, uint16_t* px, long cap,
                     char* err, int errlen) {
  try {
    nm03::dicom::Slice s = nm03::dicom::load(path);
    *cols = s.cols; *rows = s.rows; *is_signed = s.pixel_representation; *instance = s.instance_number;
    if (px) {
      if ((long)s.pixels.size() > cap) throw std::runtime_error("buffer too small");
      std::memcpy(px, s.pixels.data(), s.pixels.size() * 2);
    }
    return 0;
  } catch (const std::exception& e) { if (err) { std::strncpy(err, e.what(), errlen - 1); err[errlen - 1] = 0; } return -1; }
}

void nm03h_phantom(int w, int h, int patient, int slice, int kind, int16_t* out) {
  const std::vector<int16_t> v = nm03::phantom(w, h, patient, slice, kind ? nm03::PHANTOM_LOWCONTRAST : nm03::PHANTOM_DEFAULT);
  std::memcpy(out, v.data(), v.size() * 2);
}

long nm03h_jpeg_encode(const uint8_t* px, int w, int h, int channels, int quality, uint8_t* out, long cap) {
  try {
    const std::vector<uint8_t> b = nm03::jpeg::encode(px, w, h, channels, quality);
    if ((long)b.size() > cap) return -(long)b.size();
    std::memcpy(out, b.data(), b.size());
    return (long)b.size();
  } catch (const std::exception&) { return 0; }
}

// ImageRenderer on a 16-bit slice -> RGB canvas [ch][cw][3] over black (main_sequential.cpp:258-259)
int nm03h_render_image(const void* px, int w, int h, int is_unsigned, int cw, int ch, uint8_t* rgb) {
  try {
    auto img = std::make_shared<nm03::Image>(w, h, is_unsigned ? nm03::TYPE_UINT16 : nm03::TYPE_INT16);
    std::memcpy(img->data(), px, (size_t)w * h * 2);
    auto rti = nm03::RenderToImage::create(nm03::Color::Black(), cw, ch);
    rti->connect(nm03::ImageRenderer::create()->addInputData(img));
    rti->update();
    std::memcpy(rgb, rti->getOutputData<nm03::Image>(0)->data(), (size_t)cw * ch * 3);
    return 0;
  } catch (const std::exception&) { return -1; }
}

// SegmentationRenderer(label 1 = white, opacity, borderOpacity, borderRadius) over black (main_sequential.cpp:261)
int nm03h_render_segmentation(const uint8_t* lab, int w, int h, int cw, int ch, float opacity, float border_opacity, int radius,
                              uint8_t* rgb) {
  try {
    auto img = std::make_shared<nm03::Image>(w, h, nm03::TYPE_UINT8);
    std::memcpy(img->data(), lab, (size_t)w * h);
    nm03::LabelColors colors;
    colors[1] = nm03::Color::White();
    auto rti = nm03::RenderToImage::create(nm03::Color::Black(), cw, ch);
    rti->connect(nm03::SegmentationRenderer::create(colors, opacity, border_opacity, radius)->addInputData(img));
    rti->update();
    std::memcpy(rgb, rti->getOutputData<nm03::Image>(0)->data(), (size_t)cw * ch * 3);
    return 0;
  } catch (const std::exception&) { return -1; }
}

// the same two renders on the one-channel canvas the batch driver exports from (RenderToImage keeps an all-grey canvas as
// one plane); lab_is_bits = 1: `lab` is a bit-packed {0,1} label image (nm03_outputs.dilated_bits layout)
int nm03h_render_image_grey(const void* px, int w, int h, int is_unsigned, int cw, int ch, uint8_t* grey) {
  try {
    auto img = nm03::Image::view(px, w, h, is_unsigned ? nm03::TYPE_UINT16 : nm03::TYPE_INT16);
    auto rti = nm03::RenderToImage::create(nm03::Color::Black(), cw, ch);
    rti->connect(nm03::ImageRenderer::create()->addInputData(img));
    rti->update();
    const uint8_t* plane = rti->getOutputData<nm03::Image>(0)->greyPlane();
    if (!plane) return -2;
    std::memcpy(grey, plane, (size_t)cw * ch);
    return 0;
  } catch (const std::exception&) { return -1; }
}

int nm03h_render_segmentation_grey(const void* lab, int lab_is_bits, int w, int h, int cw, int ch, float opacity, float border_opacity,
                                   int radius, uint8_t* grey) {
  try {
    auto img = lab_is_bits ? nm03::Image::viewBits((const uint32_t*)lab, w, h) : nm03::Image::view(lab, w, h, nm03::TYPE_UINT8);
    nm03::LabelColors colors;
    colors[1] = nm03::Color::White();
    auto rti = nm03::RenderToImage::create(nm03::Color::Black(), cw, ch);
    rti->connect(nm03::SegmentationRenderer::create(colors, opacity, border_opacity, radius)->addInputData(img));
    rti->update();
    const uint8_t* plane = rti->getOutputData<nm03::Image>(0)->greyPlane();
    if (!plane) return -2;
    std::memcpy(grey, plane, (size_t)cw * ch);
    return 0;
  } catch (const std::exception&) { return -1; }
}

// CPU statement of nm03_outputs.export_blocks / export_nz: ImageRenderer's grey rendering of a 16-bit slice on a canvas
// of its own size, then level shift, transform, quantisation and zigzag per 8x8 block (w, h multiples of 8)
int nm03h_export_blocks_cpu(const void* px, int w, int h, int is_unsigned, int quality, int16_t* blocks, uint64_t* nz) {
  try {
    if (w <= 0 || h <= 0 || (w & 7) || (h & 7)) return -3;
    auto img = nm03::Image::view(px, w, h, is_unsigned ? nm03::TYPE_UINT16 : nm03::TYPE_INT16);
    auto rti = nm03::RenderToImage::create(nm03::Color::Black(), w, h);
    rti->connect(nm03::ImageRenderer::create()->addInputData(img));
    rti->update();
    const uint8_t* plane = rti->getOutputData<nm03::Image>(0)->greyPlane();
    if (!plane) return -2;
    const nm03::jpeg::Tables tb = nm03::jpeg::make_tables(quality);
    size_t b = 0;
    for (int by = 0; by < h; by += 8)
      for (int bx = 0; bx < w; bx += 8, ++b)
        nm03::jpeg::block_coefficients(plane + (size_t)by * w + bx, (size_t)w, tb.rq[0], blocks + 64 * b, nz + b);
    return 0;
  } catch (const std::exception&) { return -1; }
}

// CPU statement of nm03_outputs.export_seg_blocks: SegmentationRenderer(label 1 = white, opacity, borderOpacity, radius)
// over black on a canvas of the image's own size, then the same per-block steps; style3 receives the three greys
// (background, interior, border) a caller hands to the GPU
int nm03h_export_seg_blocks_cpu(const void* lab, int lab_is_bits, int w, int h, float opacity, float border_opacity, int radius,
                                int quality, int16_t* blocks, uint64_t* nz, uint8_t* style3) {
  try {
    if (w <= 0 || h <= 0 || (w & 7) || (h & 7)) return -3;
    auto img = lab_is_bits ? nm03::Image::viewBits((const uint32_t*)lab, w, h) : nm03::Image::view(lab, w, h, nm03::TYPE_UINT8);
    nm03::LabelColors colors;
    colors[1] = nm03::Color::White();
    auto rti = nm03::RenderToImage::create(nm03::Color::Black(), w, h);
    rti->connect(nm03::SegmentationRenderer::create(colors, opacity, border_opacity, radius)->addInputData(img));
    rti->update();
    const uint8_t* plane = rti->getOutputData<nm03::Image>(0)->greyPlane();
    if (!plane) return -2;
    const nm03::jpeg::Tables tb = nm03::jpeg::make_tables(quality);
    size_t b = 0;
    for (int by = 0; by < h; by += 8)
      for (int bx = 0; bx < w; bx += 8, ++b)
        nm03::jpeg::block_coefficients(plane + (size_t)by * w + bx, (size_t)w, tb.rq[0], blocks + 64 * b, nz + b);
    if (style3) {
      style3[0] = 0;
      style3[1] = nm03::SegmentationRenderer::blend(0, 1.0f, opacity);
      style3[2] = nm03::SegmentationRenderer::blend(0, 1.0f, border_opacity);
    }
    return 0;
  } catch (const std::exception&) { return -1; }
}

// the driver's one-step export of the result: SegmentationRenderer(label 1 = white, ...) over black + baseline JPEG, from the bit-packed mask
long nm03h_segmentation_jpeg_from_bits(const uint32_t* bits, int w, int h, float opacity, float border_opacity, int radius, int quality,
                                       uint8_t* out, long cap) {
  try {
    const std::vector<uint8_t> b = nm03::encodeSegmentationJpeg(bits, w, h, 1.0f, opacity, border_opacity, radius, 0, quality);
    if ((long)b.size() > cap) return -(long)b.size();
    std::memcpy(out, b.data(), b.size());
    return (long)b.size();
  } catch (const std::exception&) { return 0; }
}

// the reciprocal quantiser (64 floats, natural order) a caller hands to nm03_outputs.export_rq for a JPEG quality
void nm03h_jpeg_rq(int quality, float* rq64) {
  const nm03::jpeg::Tables tb = nm03::jpeg::make_tables(quality);
  std::memcpy(rq64, tb.rq[0], 64 * sizeof(float));
}

long nm03h_jpeg_from_blocks(const int16_t* blocks, const uint64_t* nz, int w, int h, int quality, uint8_t* out, long cap) {
  try {
    const std::vector<uint8_t> b = nm03::jpeg::encode_blocks(blocks, nz, w, h, quality);
    if ((long)b.size() > cap) return -(long)b.size();
    std::memcpy(out, b.data(), b.size());
    return (long)b.size();
  } catch (const std::exception&) { return 0; }
}

}  // extern "C"
