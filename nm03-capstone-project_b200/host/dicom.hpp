// dicom.hpp — minimal DICOM Part-10 reader/writer for single-frame 16-bit monochrome slices.
//
// Replaces what the reference gets from FAST's DICOMFileImporter (DCMTK underneath) on the hot path:
// `DICOMFileImporter::create(file)->setLoadSeries(false)->update()` yields ONE 2-D int16/uint16 image
// (reference src/sequential/main_sequential.cpp:175-186, src/test/test_pipeline.cpp:33-51).
// Supported: explicit and implicit VR little endian, uncompressed, BitsAllocated 16, one sample per
// pixel, nested/undefined-length sequences are skipped.  Not supported (clear error): encapsulated
// (compressed) transfer syntaxes, big endian, multi-frame.
#pragma once
#include <cstdint>
#include <cstring>
#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>

#include <fstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace nm03 {
namespace dicom {

struct Slice {
  int rows = 0, cols = 0;
  int bits_allocated = 0;
  int pixel_representation = 0;  // 0 = unsigned, 1 = two's complement
  double rescale_slope = 1.0, rescale_intercept = 0.0;
  int instance_number = 0;
  std::vector<uint16_t> pixels;  // raw stored values, row-major
};

namespace detail {

struct Cursor {
  const uint8_t* p;
  size_t n, off = 0;
  bool has(size_t k) const { return off + k <= n; }
  uint16_t u16() { if (!has(2)) throw std::runtime_error("DICOM: truncated file"); uint16_t v = (uint16_t)(p[off] | (p[off + 1] << 8)); off += 2; return v; }
  uint32_t u32() { if (!has(4)) throw std::runtime_error("DICOM: truncated file"); uint32_t v = (uint32_t)p[off] | ((uint32_t)p[off + 1] << 8) | ((uint32_t)p[off + 2] << 16) | ((uint32_t)p[off + 3] << 24); off += 4; return v; }
};

inline bool long_vr(const char* vr) {
  static const char* L[] = {"OB", "OW", "OF", "SQ", "UT", "UN", "OD", "OL", "UC", "UR", "OV", "SV", "UV"};
  for (auto s : L) if (vr[0] == s[0] && vr[1] == s[1]) return true;
  return false;
}

struct Element { uint16_t group, elem; char vr[3]; uint32_t len; size_t value_off; };

// reads one element header; value is NOT consumed
inline Element read_header(Cursor& c, bool explicit_vr) {
  Element e{};
  e.group = c.u16();
  e.elem = c.u16();
  e.vr[0] = e.vr[1] = '?'; e.vr[2] = 0;
  if (e.group == 0xFFFE) {  // item / delimiters: always implicit form
    e.len = c.u32();
  } else if (explicit_vr) {
    if (!c.has(2)) throw std::runtime_error("DICOM: truncated file");
    e.vr[0] = (char)c.p[c.off]; e.vr[1] = (char)c.p[c.off + 1];
    c.off += 2;
    if (long_vr(e.vr)) { c.u16(); e.len = c.u32(); }
    else e.len = c.u16();
  } else {
    e.len = c.u32();
  }
  e.value_off = c.off;
  return e;
}

// skip a sequence / item of undefined length by walking its content
inline void skip_undefined(Cursor& c, bool explicit_vr, bool is_item, int depth = 0) {
  if (depth > 32) throw std::runtime_error("DICOM: sequences nested deeper than 32 levels");   // untrusted input: bound the recursion
  for (;;) {
    Element e = read_header(c, explicit_vr);
    if (e.group == 0xFFFE && e.elem == 0xE0DD && !is_item) return;  // sequence delimitation
    if (e.group == 0xFFFE && e.elem == 0xE00D && is_item) return;   // item delimitation
    if (e.group == 0xFFFE && e.elem == 0xE000) {                     // item
      if (e.len == 0xFFFFFFFFu) skip_undefined(c, explicit_vr, true, depth + 1);
      else { if (!c.has(e.len)) throw std::runtime_error("DICOM: truncated item"); c.off += e.len; }
      continue;
    }
    if (e.len == 0xFFFFFFFFu) skip_undefined(c, explicit_vr, false, depth + 1);
    else { if (!c.has(e.len)) throw std::runtime_error("DICOM: truncated element"); c.off += e.len; }
  }
}

inline std::string str_value(const Cursor& c, const Element& e) {
  std::string s((const char*)c.p + e.value_off, e.len);
  while (!s.empty() && (s.back() == ' ' || s.back() == '\0')) s.pop_back();
  return s;
}

}  // namespace detail

// Header of a slice plus the location of its pixel data inside the file image (no pixel copy).
struct View {
  int rows = 0, cols = 0, bits_allocated = 0, pixel_representation = 0, instance_number = 0;
  double rescale_slope = 1.0, rescale_intercept = 0.0;
  size_t pixel_offset = 0, pixel_bytes = 0;
};

// Scan a Part-10 file image.  `data` holds the first `size` bytes of a file of `total` bytes (total = 0: the
// whole file is in memory): the header must lie inside `data`, the pixel payload only inside the file.
// Throws NeedMore if the header does not fit in `size` bytes.
struct NeedMore : std::runtime_error { NeedMore() : std::runtime_error("DICOM: header longer than the bytes read") {} };

inline View parse_view_impl(const uint8_t* data, size_t size, size_t total);

inline View parse_view(const uint8_t* data, size_t size, size_t total = 0) {
  if (total <= size) return parse_view_impl(data, size, size);
  // Only the head of the file is in memory: an element header or a nested sequence that straddles the end of the
  // buffer looks like a truncated file to the parser.  Any such error means "read more", not "bad file".
  try {
    return parse_view_impl(data, size, total);
  } catch (const NeedMore&) {
    throw;
  } catch (const std::runtime_error& e) {
    if (std::strstr(e.what(), "truncated")) throw NeedMore();
    throw;
  }
}

inline View parse_view_impl(const uint8_t* data, size_t size, size_t total) {
  using namespace detail;
  if (total < size) total = size;
  const bool partial = total > size;
  Cursor c{data, size};
  if (size < 132 || std::memcmp(data + 128, "DICM", 4) != 0) throw std::runtime_error("DICOM: missing DICM preamble");
  c.off = 132;
  std::string ts = "1.2.840.10008.1.2";
  while (c.has(8)) {
    const size_t save = c.off;
    Element e = read_header(c, true);
    if (e.group != 0x0002) { c.off = save; break; }
    if (!c.has(e.len)) throw std::runtime_error("DICOM: truncated meta header");
    if (e.elem == 0x0010) ts = str_value(c, e);
    c.off += e.len;
  }
  bool explicit_vr;
  if (ts == "1.2.840.10008.1.2") explicit_vr = false;
  else if (ts == "1.2.840.10008.1.2.1") explicit_vr = true;
  else throw std::runtime_error("DICOM: unsupported transfer syntax " + ts + " (only uncompressed little endian)");
  View s;
  int samples = 1, frames = 1;
  while (c.has(8)) {
    Element e = read_header(c, explicit_vr);
    if (e.len == 0xFFFFFFFFu) {
      if (e.group == 0x7FE0 && e.elem == 0x0010) throw std::runtime_error("DICOM: encapsulated pixel data is not supported");
      skip_undefined(c, explicit_vr, false);
      continue;
    }
    const bool is_pixels = e.group == 0x7FE0 && e.elem == 0x0010;
    if (is_pixels ? e.value_off + (size_t)e.len > total : !c.has(e.len)) {
      if (partial && !is_pixels) throw NeedMore();
      throw std::runtime_error("DICOM: truncated element");
    }
    const uint8_t* v = c.p + e.value_off;
    auto us = [&]() { return e.len >= 2 ? (int)(v[0] | (v[1] << 8)) : 0; };
    if (e.group == 0x0028) {
      switch (e.elem) {
        case 0x0002: samples = us(); break;
        case 0x0008: frames = std::atoi(str_value(c, e).c_str()); break;
        case 0x0010: s.rows = us(); break;
        case 0x0011: s.cols = us(); break;
        case 0x0100: s.bits_allocated = us(); break;
        case 0x0103: s.pixel_representation = us(); break;
        case 0x1052: s.rescale_intercept = std::atof(str_value(c, e).c_str()); break;
        case 0x1053: s.rescale_slope = std::atof(str_value(c, e).c_str()); break;
        default: break;
      }
    } else if (e.group == 0x0020 && e.elem == 0x0013) {
      s.instance_number = std::atoi(str_value(c, e).c_str());
    } else if (e.group == 0x7FE0 && e.elem == 0x0010) {
      if (s.rows <= 0 || s.cols <= 0) throw std::runtime_error("DICOM: pixel data before image size");
      if (s.bits_allocated != 16) throw std::runtime_error("DICOM: only 16-bit pixels are supported");
      if (samples != 1) throw std::runtime_error("DICOM: only one sample per pixel is supported");
      if (frames > 1) throw std::runtime_error("DICOM: multi-frame images are not supported");
      const size_t need = (size_t)s.rows * s.cols * 2;
      if (e.len < need) throw std::runtime_error("DICOM: pixel data shorter than rows*cols");
      s.pixel_offset = e.value_off;
      s.pixel_bytes = need;
      return s;
    }
    c.off += e.len;
  }
  if (partial) throw NeedMore();
  throw std::runtime_error("DICOM: no pixel data element");
}

// Modality LUT (PS3.3 C.11.1): value = RescaleSlope * stored + RescaleIntercept.  FAST's importer hands the pipeline
// the pixel values DCMTK reports (src/sequential/main_sequential.cpp:175-186), which include this step, so it is
// applied here, in place, before the slice goes to the GPU.  The pipeline's input is a 16-bit integer image: slope and
// intercept must be integers and every rescaled value must fit the output type — signed 16 bit when the stored values
// are signed or the intercept is negative, unsigned 16 bit otherwise.  Anything else is refused with a clear error
// rather than segmented from the wrong values.  Returns the pixel representation of the result (0 unsigned, 1 signed).
inline int rescaled_representation(int pixel_representation, double slope, double intercept) {
  if (slope == 1.0 && intercept == 0.0) return pixel_representation;
  return (pixel_representation == 1 || intercept < 0.0) ? 1 : 0;
}

inline int rescale_in_place(uint16_t* px, size_t n, int pixel_representation, double slope, double intercept) {
  if (slope == 1.0 && intercept == 0.0) return pixel_representation;
  if (slope != (double)(long long)slope || intercept != (double)(long long)intercept || slope == 0.0)
    throw std::runtime_error("DICOM: RescaleSlope " + std::to_string(slope) + " / RescaleIntercept " + std::to_string(intercept) +
                             " is not supported (integer slope and intercept only)");
  const long long a = (long long)slope, b = (long long)intercept;
  const int out_rep = rescaled_representation(pixel_representation, slope, intercept);
  const long long lo = out_rep ? -32768 : 0, hi = out_rep ? 32767 : 65535;
  for (size_t i = 0; i < n; ++i) {
    const long long stored = pixel_representation ? (long long)(int16_t)px[i] : (long long)px[i];
    const long long v = a * stored + b;
    if (v < lo || v > hi) throw std::runtime_error("DICOM: rescaled pixel value " + std::to_string(v) + " does not fit 16 bits");
    px[i] = (uint16_t)(int64_t)v;
  }
  return out_rep;
}

// Parse a Part-10 file image held in memory (header + a copy of the pixels; the modality LUT is applied).
inline Slice parse(const uint8_t* data, size_t size) {
  const View v = parse_view(data, size);
  Slice s;
  s.rows = v.rows; s.cols = v.cols; s.bits_allocated = v.bits_allocated; s.pixel_representation = v.pixel_representation;
  s.rescale_slope = v.rescale_slope; s.rescale_intercept = v.rescale_intercept; s.instance_number = v.instance_number;
  s.pixels.resize((size_t)v.rows * v.cols);
  std::memcpy(s.pixels.data(), data + v.pixel_offset, v.pixel_bytes);  // little endian host assumed (x86-64 / aarch64)
  s.pixel_representation = rescale_in_place(s.pixels.data(), s.pixels.size(), v.pixel_representation, v.rescale_slope, v.rescale_intercept);
  return s;
}

// Positional reads of one file: the feeder reads the first bytes, parses the header there, and then
// reads the pixel payload straight into pinned staging memory (one copy, kernel -> pinned).
class FileReader {
 public:
  FileReader() = default;
  explicit FileReader(const std::string& path) {
    fd_ = ::open(path.c_str(), O_RDONLY);
    if (fd_ < 0) throw std::runtime_error("cannot open " + path);
    struct stat st;
    if (::fstat(fd_, &st) != 0 || st.st_size <= 0) { ::close(fd_); fd_ = -1; throw std::runtime_error("cannot stat " + path); }
    size_ = (size_t)st.st_size;
  }
  FileReader(FileReader&& o) noexcept : size_(o.size_), fd_(o.fd_) { o.size_ = 0; o.fd_ = -1; }
  FileReader& operator=(FileReader&& o) noexcept {
    if (this != &o) { if (fd_ >= 0) ::close(fd_); size_ = o.size_; fd_ = o.fd_; o.size_ = 0; o.fd_ = -1; }
    return *this;
  }
  FileReader(const FileReader&) = delete;
  FileReader& operator=(const FileReader&) = delete;
  ~FileReader() { if (fd_ >= 0) ::close(fd_); }
  size_t size() const { return size_; }
  void read_at(void* dst, size_t bytes, size_t offset) const {
    uint8_t* d = (uint8_t*)dst;
    while (bytes > 0) {
      const ssize_t n = ::pread(fd_, d, bytes, (off_t)offset);
      if (n <= 0) throw std::runtime_error("DICOM: short read");
      d += n; offset += (size_t)n; bytes -= (size_t)n;
    }
  }
  // header of the slice; reads 16 KiB first and the whole file only if the header is longer than that
  View view() const {
    std::vector<uint8_t> head(std::min<size_t>(size_, 16384));
    read_at(head.data(), head.size(), 0);
    try {
      return parse_view(head.data(), head.size(), size_);
    } catch (const NeedMore&) {
      head.resize(size_);
      read_at(head.data(), size_, 0);
      return parse_view(head.data(), size_, size_);
    }
  }
 private:
  size_t size_ = 0;
  int fd_ = -1;
};

inline std::vector<uint8_t> read_file(const std::string& path) {
  std::ifstream f(path, std::ios::binary);
  if (!f) throw std::runtime_error("cannot open " + path);
  f.seekg(0, std::ios::end);
  const std::streamsize n = f.tellg();
  f.seekg(0);
  std::vector<uint8_t> buf((size_t)n);
  if (n > 0 && !f.read((char*)buf.data(), n)) throw std::runtime_error("cannot read " + path);
  return buf;
}

inline Slice load(const std::string& path) {
  const std::vector<uint8_t> buf = read_file(path);
  return parse(buf.data(), buf.size());
}

// ---- writer (synthetic inputs: "identical synthetic DICOM inputs") -------------------
namespace detail {
struct Out {
  std::vector<uint8_t> b;
  void u16(uint16_t v) { b.push_back((uint8_t)v); b.push_back((uint8_t)(v >> 8)); }
  void u32(uint32_t v) { for (int i = 0; i < 4; ++i) b.push_back((uint8_t)(v >> (8 * i))); }
  void bytes(const void* p, size_t n) { const uint8_t* q = (const uint8_t*)p; b.insert(b.end(), q, q + n); }
  void elem(bool explicit_vr, uint16_t g, uint16_t e, const char* vr, const void* val, uint32_t len) {
    u16(g); u16(e);
    if (explicit_vr) {
      b.push_back((uint8_t)vr[0]); b.push_back((uint8_t)vr[1]);
      if (long_vr(vr)) { u16(0); u32(len); } else u16((uint16_t)len);
    } else {
      u32(len);
    }
    bytes(val, len);
  }
  void str(bool ex, uint16_t g, uint16_t e, const char* vr, std::string s, char pad = ' ') {
    if (s.size() & 1) s.push_back(pad);
    elem(ex, g, e, vr, s.data(), (uint32_t)s.size());
  }
  void us(bool ex, uint16_t g, uint16_t e, uint16_t v) { uint8_t t[2] = {(uint8_t)v, (uint8_t)(v >> 8)}; elem(ex, g, e, "US", t, 2); }
};
}  // namespace detail

inline std::vector<uint8_t> encode(const uint16_t* pixels, int cols, int rows, bool is_signed, int instance_number,
                                   bool explicit_vr = true, const std::string& slope = "1", const std::string& intercept = "0") {
  using detail::Out;
  const std::string ts = explicit_vr ? "1.2.840.10008.1.2.1" : "1.2.840.10008.1.2";
  Out meta;
  const uint8_t ver[2] = {0, 1};
  meta.elem(true, 0x0002, 0x0001, "OB", ver, 2);
  meta.str(true, 0x0002, 0x0002, "UI", "1.2.840.10008.5.1.4.1.1.4", '\0');  // MR Image Storage
  meta.str(true, 0x0002, 0x0003, "UI", "1.2.826.0.1.3680043.9.7433.1." + std::to_string(instance_number), '\0');
  meta.str(true, 0x0002, 0x0010, "UI", ts, '\0');
  meta.str(true, 0x0002, 0x0012, "UI", "1.2.826.0.1.3680043.9.7433.0.1", '\0');
  Out o;
  o.b.assign(128, 0);
  o.bytes("DICM", 4);
  const uint32_t glen = (uint32_t)meta.b.size();
  uint8_t gl[4] = {(uint8_t)glen, (uint8_t)(glen >> 8), (uint8_t)(glen >> 16), (uint8_t)(glen >> 24)};
  o.elem(true, 0x0002, 0x0000, "UL", gl, 4);
  o.bytes(meta.b.data(), meta.b.size());
  const bool ex = explicit_vr;
  o.str(ex, 0x0008, 0x0060, "CS", "MR");
  o.str(ex, 0x0020, 0x0013, "IS", std::to_string(instance_number));
  o.us(ex, 0x0028, 0x0002, 1);
  o.str(ex, 0x0028, 0x0004, "CS", "MONOCHROME2");
  o.us(ex, 0x0028, 0x0010, (uint16_t)rows);
  o.us(ex, 0x0028, 0x0011, (uint16_t)cols);
  o.us(ex, 0x0028, 0x0100, 16);
  o.us(ex, 0x0028, 0x0101, 16);
  o.us(ex, 0x0028, 0x0102, 15);
  o.us(ex, 0x0028, 0x0103, is_signed ? 1 : 0);
  o.str(ex, 0x0028, 0x1052, "DS", intercept);
  o.str(ex, 0x0028, 0x1053, "DS", slope);
  o.elem(ex, 0x7FE0, 0x0010, "OW", pixels, (uint32_t)((size_t)rows * cols * 2));
  return o.b;
}

inline void save(const std::string& path, const uint16_t* pixels, int cols, int rows, bool is_signed, int instance_number,
                 bool explicit_vr = true, const std::string& slope = "1", const std::string& intercept = "0") {
  const std::vector<uint8_t> b = encode(pixels, cols, rows, is_signed, instance_number, explicit_vr, slope, intercept);
  std::ofstream f(path, std::ios::binary);
  if (!f || !f.write((const char*)b.data(), (std::streamsize)b.size())) throw std::runtime_error("cannot write " + path);
}

}  // namespace dicom
}  // namespace nm03
