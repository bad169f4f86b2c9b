// host_capi.cpp — C entry points over the host-side helpers (DICOM reader/writer, phantom, JPEG,
// renderer restatement) so that they can be exercised from tests without a GPU (libnm03_host.so).
#include <cstring>

#include "dicom.hpp"
#include "jpeg.hpp"
#include "phantom.hpp"

extern "C" {

// returns 0 or -1 (message in err)
int nm03h_dicom_write(const char* path, const uint16_t* px, int cols, int rows, int is_signed, int instance, int explicit_vr,
                      char* err, int errlen) {
  try { nm03::dicom::save(path, px, cols, rows, is_signed != 0, instance, explicit_vr != 0); return 0; }
  catch (const std::exception& e) { if (err) std::strncpy(err, e.what(), errlen - 1); return -1; }
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

}  // extern "C"
