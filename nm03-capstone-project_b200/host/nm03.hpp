// nm03.hpp — FAST-shaped C++17 shim over the C ABI (include/nm03.h).
//
// The reference drives FAST through `T::create(..)->connect(..)->update()` and
// `getOutputData<Image>(0)` (src/sequential/main_sequential.cpp:175-262,
// src/parallel/main_parallel.cpp:78-210, src/test/test_pipeline.cpp:33-177).  This header declares
// classes with those names, members and error behaviour so that the body of processSingleImage()
// compiles unchanged (`namespace fast = nm03;`, see host/compat/FAST/FAST_directives.hpp), while the
// arithmetic runs in libnm03.so on the GPU.
//
// Evaluation is lazy and fused: update() on the preprocessing / segmentation stages only records the
// request; Dilation::update() / Erosion::update() (or the first getOutputData() that needs pixels)
// recognises the reference's chain  importer -> normalise -> clip -> median -> sharpen -> region
// growing -> caster -> dilation|erosion  and issues ONE nm03_process_batch call for the slice.
// Anything FAST would throw surfaces as nm03::Exception from that call, inside the same try block
// the reference uses (main_sequential.cpp:267, main_parallel.cpp:163).
//
// Rendering (ImageRenderer / SegmentationRenderer / RenderToImage) is FAST's OpenGL path in the
// reference; a graphics stack is unavailable on this box, so it is restated on the CPU (window/level,
// label overlay, aspect-fit into the canvas).  Pixel parity with FAST's renderer is not claimed.
#pragma once
#include <nm03.h>
#if defined(__SSE2__)
#include <emmintrin.h>
#endif

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "dicom.hpp"
#include "jpeg.hpp"

namespace nm03 {

class Exception : public std::exception {
  std::string msg_;
 public:
  explicit Exception(std::string m) : msg_(std::move(m)) {}
  const char* what() const noexcept override { return msg_.c_str(); }
};

struct Vector3i {
  int v[3];
  Vector3i(int x = 0, int y = 0, int z = 0) : v{x, y, z} {}
  int x() const { return v[0]; }
  int y() const { return v[1]; }
  int z() const { return v[2]; }
  int operator[](int i) const { return v[i]; }
};

enum DataType { TYPE_FLOAT, TYPE_UINT8, TYPE_INT8, TYPE_UINT16, TYPE_INT16 };

struct Color {
  float r, g, b;
  static Color White() { return {1, 1, 1}; }
  static Color Black() { return {0, 0, 0}; }
  static Color Red() { return {1, 0, 0}; }
  static Color Green() { return {0, 1, 0}; }
};
using LabelColors = std::map<int, Color>;

struct Config {
  // FAST's Config::getTestDataPath() (main_sequential.cpp:83): here the NM03_TEST_DATA_PATH variable
  static std::string getTestDataPath() {
    const char* e = std::getenv("NM03_TEST_DATA_PATH");
    std::string p = e ? e : "./data/";
    if (!p.empty() && p.back() != '/') p.push_back('/');
    return p;
  }
};

struct Reporter {  // logging configuration is a no-op (main_sequential.cpp:310-315)
  enum Type { INFO, WARNING, ERROR };
  enum Method { NONE, COUT };
  static void setGlobalReportMethod(Type, Method) {}
};

struct QApplication {  // main_parallel.cpp:391 needs one only for FAST's GL context
  QApplication(int&, char**) {}
};

// ---- data objects ----------------------------------------------------------------------------
class Image {
 public:
  Image(int w, int h, DataType t, int channels = 1) : w_(w), h_(h), c_(channels), type_(t) {
    bytes_.resize((size_t)w * h * channels * elemSize(t));
  }
  // a view of pixels owned by someone else (the batch driver's pinned staging buffers): no copy, caller keeps them alive
  static std::shared_ptr<Image> view(const void* pixels, int w, int h, DataType t) {
    std::shared_ptr<Image> im(new Image(w, h, t, 1, nullptr));
    im->ext_ = pixels;
    return im;
  }
  // a {0,1} label image held bit-packed (nm03_outputs.dilated_bits: bit x%32 of word x/32 of a row = pixel x), owned by
  // someone else; the bytes are only unpacked if somebody asks for them (the segmentation renderer works on the words)
  static std::shared_ptr<Image> viewBits(const uint32_t* bits, int w, int h) {
    std::shared_ptr<Image> im(new Image(w, h, TYPE_UINT8, 1, nullptr));
    im->bits_ = bits;
    return im;
  }
  const uint32_t* labelBits() const { return bits_; }
  // an RGB canvas all of whose pixels are grey, kept as ONE channel until somebody asks for the interleaved bytes
  static std::shared_ptr<Image> greyCanvas(int w, int h, std::vector<uint8_t> plane) {
    std::shared_ptr<Image> im(new Image(w, h, TYPE_UINT8, 3, nullptr));
    im->grey_ = std::move(plane);
    return im;
  }
  int getWidth() const { return w_; }
  int getHeight() const { return h_; }
  int getNrOfChannels() const { return c_; }
  DataType getDataType() const { return type_; }
  void* data() { return const_cast<void*>(bytes()); }
  const void* data() const { return bytes(); }
  template <class T> T* as() { return reinterpret_cast<T*>(const_cast<void*>(bytes())); }
  template <class T> const T* as() const { return reinterpret_cast<const T*>(bytes()); }
  static size_t elemSize(DataType t) { return t == TYPE_FLOAT ? 4 : (t == TYPE_UINT8 || t == TYPE_INT8) ? 1 : 2; }
  // the single-channel form of a grey canvas (null for any other image); the JPEG writer takes it as it is
  const uint8_t* greyPlane() const { return grey_.empty() ? nullptr : grey_.data(); }
 private:
  Image(int w, int h, DataType t, int channels, std::nullptr_t) : w_(w), h_(h), c_(channels), type_(t) {}
  const void* bytes() const {
    if (ext_) return ext_;
    if (bytes_.empty() && bits_) {            // first access to the bytes of a bit-packed label image
      bytes_.resize((size_t)w_ * h_);
      const int wpr = (w_ + 31) / 32;
      for (int y = 0; y < h_; ++y)
        for (int x = 0; x < w_; ++x) bytes_[(size_t)y * w_ + x] = (bits_[(size_t)y * wpr + (x >> 5)] >> (x & 31)) & 1u;
    }
    if (bytes_.empty() && !grey_.empty()) {   // first access to the interleaved form of a grey canvas
      bytes_.resize(grey_.size() * 3);
      for (size_t i = 0; i < grey_.size(); ++i) bytes_[3 * i] = bytes_[3 * i + 1] = bytes_[3 * i + 2] = grey_[i];
    }
    return bytes_.data();
  }
  int w_, h_, c_;
  DataType type_;
  mutable std::vector<uint8_t> bytes_;
  std::vector<uint8_t> grey_;
  const void* ext_ = nullptr;
  const uint32_t* bits_ = nullptr;
};
using Segmentation = Image;

// ---- runtime: one context over the visible GPUs, slices assigned round-robin per host thread ----
class Runtime {
 public:
  static Runtime& get() {
    static Runtime rt;
    return rt;
  }
  nm03_ctx* ctx() const { return ctx_; }
  int gpuForThisThread() {
    thread_local int g = -1;
    if (g < 0) g = next_.fetch_add(1) % ngpus_;
    return g;
  }
  static void check(int rc) {
    if (rc != NM03_OK) throw Exception(nm03_last_error());
  }
 private:
  Runtime() {
    std::vector<int> ids;
    if (const char* e = std::getenv("NM03_GPUS")) {
      for (const char* p = e; *p;) { ids.push_back(std::atoi(p)); while (*p && *p != ',') ++p; if (*p) ++p; }
    }
    int rc = nm03_init(ids.empty() ? nullptr : ids.data(), (int)ids.size(), &ctx_);
    if (rc != NM03_OK) throw Exception(nm03_last_error());
    ngpus_ = nm03_gpu_count(ctx_);
  }
  ~Runtime() { nm03_shutdown(ctx_); }
  nm03_ctx* ctx_ = nullptr;
  int ngpus_ = 1;
  std::atomic<int> next_{0};
};

// ---- process objects ---------------------------------------------------------------------------
class ProcessObject : public std::enable_shared_from_this<ProcessObject> {
 public:
  enum Kind { IMPORTER, NORMALIZE, CLIP, MEDIAN, SHARPEN, REGION_GROW, CASTER, DILATE, ERODE, RENDERER, RENDER_TO_IMAGE, EXPORTER, OTHER };
  virtual ~ProcessObject() = default;
  virtual void update() {}
  template <class T = Image>
  std::shared_ptr<T> getOutputData(int /*port*/ = 0) {
    if (!output_) evaluate();
    return output_;
  }
  Kind kind() const { return kind_; }
  std::shared_ptr<ProcessObject> parent() const { return parent_; }
  // input image of this object: the connected data object, or the parent's output
  std::shared_ptr<Image> inputImage() {
    if (inputData_) return inputData_;
    if (!parent_) throw Exception("process object has no input connection");
    return parent_->getOutputData<Image>(0);
  }
  bool hasOutput() const { return (bool)output_; }
  void setOutput(std::shared_ptr<Image> img) { output_ = std::move(img); }
 protected:
  explicit ProcessObject(Kind k) : kind_(k) {}
  virtual void evaluate() { throw Exception("this process object produces no image"); }
  void connectParent(std::shared_ptr<ProcessObject> p) { parent_ = std::move(p); output_.reset(); }
  void connectData(std::shared_ptr<Image> d) { inputData_ = std::move(d); output_.reset(); }
  Kind kind_;
  std::shared_ptr<ProcessObject> parent_;
  std::shared_ptr<Image> inputData_;
  std::shared_ptr<Image> output_;
};

// typed create()/connect() as FAST's FAST_PROCESS_OBJECT/FAST_CONSTRUCTOR macros generate them
template <class D>
class Typed : public ProcessObject {
 public:
  template <class... A>
  static std::shared_ptr<D> create(A&&... a) { return std::shared_ptr<D>(new D(std::forward<A>(a)...)); }
  std::shared_ptr<D> connect(std::shared_ptr<ProcessObject> p) {
    connectParent(std::move(p));
    return self();
  }
  std::shared_ptr<D> connect(std::shared_ptr<Image> d) {
    connectData(std::move(d));
    return self();
  }
  std::shared_ptr<D> addInputData(std::shared_ptr<Image> d) { return connect(std::move(d)); }
 protected:
  explicit Typed(Kind k) : ProcessObject(k) {}
  std::shared_ptr<D> self() { return std::static_pointer_cast<D>(shared_from_this()); }
};

class DICOMFileImporter : public Typed<DICOMFileImporter> {
 public:
  void setLoadSeries(bool) {}  // a single 2-D slice is all this importer ever loads (test_pipeline.cpp:38-42)
  void update() override { if (!output_) evaluate(); }
  bool isUnsigned() const { return unsigned_; }
 protected:
  friend class Typed<DICOMFileImporter>;
  explicit DICOMFileImporter(std::string file) : Typed(IMPORTER), file_(std::move(file)) {}
  void evaluate() override {
    try {
      dicom::Slice s = dicom::load(file_);
      unsigned_ = s.pixel_representation == 0;
      auto img = std::make_shared<Image>(s.cols, s.rows, unsigned_ ? TYPE_UINT16 : TYPE_INT16);
      std::memcpy(img->data(), s.pixels.data(), s.pixels.size() * 2);
      output_ = img;
    } catch (const std::exception& e) {
      throw Exception(e.what());
    }
  }
  std::string file_;
  bool unsigned_ = false;
};

class IntensityNormalization : public Typed<IntensityNormalization> {
 public:
  float low, high, mn, mx;
 protected:
  friend class Typed<IntensityNormalization>;
  IntensityNormalization(float valueLow = 0.0f, float valueHigh = 1.0f, float minIntensity = NAN, float maxIntensity = NAN)
      : Typed(NORMALIZE), low(valueLow), high(valueHigh), mn(minIntensity), mx(maxIntensity) {}
  void evaluate() override { throw Exception("IntensityNormalization output is only available fused with median + sharpen"); }
};

class IntensityClipping : public Typed<IntensityClipping> {
 public:
  float lo, hi;
 protected:
  friend class Typed<IntensityClipping>;
  IntensityClipping(float minValue = 0.0f, float maxValue = 4000.0f) : Typed(CLIP), lo(minValue), hi(maxValue) {}
  void evaluate() override { throw Exception("IntensityClipping output is only available fused with median + sharpen"); }
};

struct Chain;  // forward

class VectorMedianFilter : public Typed<VectorMedianFilter> {
 public:
  int window;
 protected:
  friend class Typed<VectorMedianFilter>;
  explicit VectorMedianFilter(int size = 3) : Typed(MEDIAN), window(size) {}
  void evaluate() override;
};

class ImageSharpening : public Typed<ImageSharpening> {
 public:
  float gain, sigma;
  int mask;
 protected:
  friend class Typed<ImageSharpening>;
  ImageSharpening(float g = 1.0f, float stdDev = 0.5f, int maskSize = 3) : Typed(SHARPEN), gain(g), sigma(stdDev), mask(maskSize) {}
  void evaluate() override;
};

class SeededRegionGrowing : public Typed<SeededRegionGrowing> {
 public:
  float lo, hi;
  std::vector<int> seeds;  // x,y pairs
  void addSeedPoint(int x, int y) { seeds.push_back(x); seeds.push_back(y); output_.reset(); }
  void addSeedPoint(Vector3i p) { addSeedPoint(p.x(), p.y()); }
 protected:
  friend class Typed<SeededRegionGrowing>;
  SeededRegionGrowing(float intensityMinimum, float intensityMaximum, std::vector<Vector3i> seedPoints = {})
      : Typed(REGION_GROW), lo(intensityMinimum), hi(intensityMaximum) {
    for (auto& s : seedPoints) addSeedPoint(s);
  }
  void evaluate() override;
};

class ImageCaster : public Typed<ImageCaster> {
 public:
  DataType type;
  float scale;
 protected:
  friend class Typed<ImageCaster>;
  explicit ImageCaster(DataType outputType, float scaleFactor = 1.0f) : Typed(CASTER), type(outputType), scale(scaleFactor) {}
  void evaluate() override {
    auto in = inputImage();
    if (type != TYPE_UINT8 || in->getDataType() != TYPE_UINT8)
      throw Exception("ImageCaster: only the uint8 -> uint8 cast of the reference pipeline is implemented");
    output_ = in;  // identity on a {0,1} segmentation (main_sequential.cpp:246-248)
  }
};

class Morphology : public ProcessObject {
 public:
  int size;
  void update() override { if (!output_) evaluate(); }
 protected:
  Morphology(Kind k, int s) : ProcessObject(k), size(s) {}
  void evaluate() override;
};

class Dilation : public Morphology {
 public:
  static std::shared_ptr<Dilation> create(int size = 3) { return std::shared_ptr<Dilation>(new Dilation(size)); }
  std::shared_ptr<Dilation> connect(std::shared_ptr<ProcessObject> p) { connectParent(std::move(p)); return std::static_pointer_cast<Dilation>(shared_from_this()); }
 private:
  explicit Dilation(int s) : Morphology(DILATE, s) {}
};

class Erosion : public Morphology {
 public:
  static std::shared_ptr<Erosion> create(int size = 3) { return std::shared_ptr<Erosion>(new Erosion(size)); }
  std::shared_ptr<Erosion> connect(std::shared_ptr<ProcessObject> p) { connectParent(std::move(p)); return std::static_pointer_cast<Erosion>(shared_from_this()); }
 private:
  explicit Erosion(int s) : Morphology(ERODE, s) {}
};

// ---- chain recognition + fused evaluation ---------------------------------------------------------
struct Chain {
  std::shared_ptr<Image> raw;
  IntensityNormalization* norm = nullptr;
  IntensityClipping* clip = nullptr;
  VectorMedianFilter* median = nullptr;
  ImageSharpening* sharpen = nullptr;
  SeededRegionGrowing* rg = nullptr;
  nm03_params params;

  // walk up from `top` (a MEDIAN, SHARPEN or REGION_GROW node)
  static Chain fromNode(ProcessObject* top) {
    Chain c;
    nm03_default_params(&c.params);
    ProcessObject* n = top;
    if (n->kind() == ProcessObject::REGION_GROW) { c.rg = static_cast<SeededRegionGrowing*>(n); n = n->parent().get(); }
    if (n && n->kind() == ProcessObject::SHARPEN) { c.sharpen = static_cast<ImageSharpening*>(n); n = n->parent().get(); }
    if (n && n->kind() == ProcessObject::MEDIAN) { c.median = static_cast<VectorMedianFilter*>(n); n = n->parent().get(); }
    if (n && n->kind() == ProcessObject::CLIP) { c.clip = static_cast<IntensityClipping*>(n); n = n->parent().get(); }
    if (n && n->kind() == ProcessObject::NORMALIZE) { c.norm = static_cast<IntensityNormalization*>(n); n = n->parent().get(); }
    if (!n || !c.norm || !c.clip || !c.median)
      throw Exception("unsupported operator chain: expected importer -> IntensityNormalization -> IntensityClipping -> "
                      "VectorMedianFilter -> ImageSharpening [-> SeededRegionGrowing]");
    if (std::isnan(c.norm->mn) || std::isnan(c.norm->mx))
      throw Exception("IntensityNormalization without explicit min/max intensity is not implemented");
    c.raw = n->getOutputData<Image>(0);
    if (!c.raw) throw Exception("Failed to get imported image");
    if (c.raw->getDataType() != TYPE_INT16 && c.raw->getDataType() != TYPE_UINT16)
      throw Exception("pipeline input must be a 16-bit image");
    nm03_params& p = c.params;
    p.norm_low = c.norm->low; p.norm_high = c.norm->high; p.norm_min = c.norm->mn; p.norm_max = c.norm->mx;
    p.clip_min = c.clip->lo; p.clip_max = c.clip->hi;
    p.median_window = c.median->window;
    if (c.sharpen) { p.sharpen_gain = c.sharpen->gain; p.sharpen_sigma = c.sharpen->sigma; p.sharpen_mask = c.sharpen->mask; }
    if (c.rg) { p.rg_min = c.rg->lo; p.rg_max = c.rg->hi; }
    p.input_unsigned = c.raw->getDataType() == TYPE_UINT16;
    return c;
  }

  void preprocess(bool want_sharpened) {
    Runtime& rt = Runtime::get();
    const int w = raw->getWidth(), h = raw->getHeight();
    auto med = std::make_shared<Image>(w, h, TYPE_FLOAT);
    std::shared_ptr<Image> sh = want_sharpened ? std::make_shared<Image>(w, h, TYPE_FLOAT) : nullptr;
    Runtime::check(nm03_preprocess(rt.ctx(), rt.gpuForThisThread(), raw->data(), w, h, &params, med->as<float>(),
                                   sh ? sh->as<float>() : nullptr, nullptr));
    median->setOutput(med);
    if (sh) sharpen->setOutput(sh);
  }

  // whole pipeline for the slice in one call; morph_size 0 = stop after region growing
  void pipeline(int dilate_size, std::shared_ptr<Image>* dilated, int erode_size, std::shared_ptr<Image>* eroded) {
    Runtime& rt = Runtime::get();
    const int w = raw->getWidth(), h = raw->getHeight();
    if (!sharpen) throw Exception("SeededRegionGrowing must be connected to ImageSharpening in the fused pipeline");
    auto mask = std::make_shared<Image>(w, h, TYPE_UINT8);
    nm03_outputs out = {};
    out.rg_mask = mask->as<uint8_t>();
    // FAST validates Dilation/Erosion sizes even when only one of them is requested here
    params.dilate_size = dilate_size > 0 ? dilate_size : 3;
    params.erode_size = erode_size;
    if (dilate_size > 0) { *dilated = std::make_shared<Image>(w, h, TYPE_UINT8); out.dilated = (*dilated)->as<uint8_t>(); }
    if (erode_size > 0) { *eroded = std::make_shared<Image>(w, h, TYPE_UINT8); out.eroded = (*eroded)->as<uint8_t>(); }
    Runtime::check(nm03_process_batch(rt.ctx(), rt.gpuForThisThread(), raw->data(), 1, w, h, &params,
                                      rg->seeds.empty() ? nullptr : rg->seeds.data(), (int)rg->seeds.size() / 2, &out, nullptr));
    rg->setOutput(mask);
  }
};

inline void VectorMedianFilter::evaluate() {
  Chain c = Chain::fromNode(this);
  c.preprocess(false);
}

inline void ImageSharpening::evaluate() {
  Chain c = Chain::fromNode(this);
  c.preprocess(true);
}

inline void SeededRegionGrowing::evaluate() {
  if (seeds.empty()) throw Exception("SeededRegionGrowing: no seed points");
  if (parent_ && parent_->hasOutput()) {  // stage-by-stage: the input image already exists
    auto in = parent_->getOutputData<Image>(0);
    if (in->getDataType() != TYPE_FLOAT) throw Exception("SeededRegionGrowing: only float input is implemented");
    nm03_params p;
    nm03_default_params(&p);
    p.rg_min = lo; p.rg_max = hi;
    auto mask = std::make_shared<Image>(in->getWidth(), in->getHeight(), TYPE_UINT8);
    Runtime& rt = Runtime::get();
    Runtime::check(nm03_region_grow(rt.ctx(), rt.gpuForThisThread(), in->as<float>(), in->getWidth(), in->getHeight(), &p,
                                    seeds.data(), (int)seeds.size() / 2, mask->as<uint8_t>(), nullptr));
    output_ = mask;
    return;
  }
  Chain c = Chain::fromNode(this);
  c.pipeline(0, nullptr, 0, nullptr);
}

inline void Morphology::evaluate() {
  if (!parent_) throw Exception("morphology operator has no input connection");
  // the reference's chain: (Dilation|Erosion) <- ImageCaster <- SeededRegionGrowing <- ... nothing computed yet
  ProcessObject* n = parent_.get();
  ProcessObject* caster = nullptr;
  if (n->kind() == CASTER) { caster = n; n = n->parent().get(); }
  if (n && n->kind() == REGION_GROW && !n->hasOutput() && !(n->parent() && n->parent()->hasOutput())) {
    Chain c = Chain::fromNode(n);
    std::shared_ptr<Image> res;
    if (kind_ == DILATE) c.pipeline(size, &res, 0, nullptr);
    else {
      if (size <= 2 || (size & 1) == 0) throw Exception("morphology size must be odd and > 2");
      c.pipeline(0, nullptr, size, &res);
    }
    if (caster) caster->setOutput(n->getOutputData<Image>(0));
    output_ = res;
    return;
  }
  auto in = parent_->getOutputData<Image>(0);
  if (in->getDataType() != TYPE_UINT8) throw Exception("Dilation/Erosion need a TYPE_UINT8 image (cast it first)");
  auto out = std::make_shared<Image>(in->getWidth(), in->getHeight(), TYPE_UINT8);
  Runtime& rt = Runtime::get();
  Runtime::check(nm03_morphology(rt.ctx(), rt.gpuForThisThread(), in->as<uint8_t>(), in->getWidth(), in->getHeight(), size,
                                 NM03_SE_DISC, kind_ == DILATE ? NM03_MORPH_DILATE : NM03_MORPH_ERODE, out->as<uint8_t>()));
  output_ = out;
}

// ---- rendering + export (CPU restatement of FAST's GL renderers; host side, e2e figure only) ------
class Renderer : public ProcessObject {
 public:
  // paint this renderer's input into a canvas (aspect-fit, centred): stride 3 = interleaved RGB, stride 1 = one grey
  // channel (only asked of renderers whose paintsGrey() is true)
  virtual void paint(uint8_t* rgb, int stride, int cw, int ch) = 0;
  virtual bool paintsGrey() const { return false; }   // true: only ever writes r == g == b
 protected:
  Renderer() : ProcessObject(RENDERER) {}
  struct Fit { float scale; float ox, oy; };
  static Fit fit(int iw, int ih, int cw, int ch) {
    const float s = std::min((float)cw / iw, (float)ch / ih);
    return {s, (cw - iw * s) * 0.5f, (ch - ih * s) * 0.5f};
  }
};

class ImageRenderer : public Renderer {
 public:
  static std::shared_ptr<ImageRenderer> create() { return std::shared_ptr<ImageRenderer>(new ImageRenderer()); }
  std::shared_ptr<ImageRenderer> connect(std::shared_ptr<ProcessObject> p) { connectParent(std::move(p)); return me(); }
  std::shared_ptr<ImageRenderer> addInputData(std::shared_ptr<Image> d) { connectData(std::move(d)); return me(); }
  bool paintsGrey() const override { return true; }
  void paint(uint8_t* rgb, int stride, int cw, int ch) override {
    auto img = inputImage();
    const int w = img->getWidth(), h = img->getHeight();
    const size_t n = (size_t)w * h;
    if (w == cw && h == ch && (img->getDataType() == TYPE_INT16 || img->getDataType() == TYPE_UINT16)) {
      // the driver's case (16-bit slice on a canvas of its own size): same window and rounding as below, but the grey
      // level of every intensity between min and max is tabulated once instead of being computed per pixel
      const bool sg = img->getDataType() == TYPE_INT16;
      const int16_t* ss = img->as<int16_t>();
      const uint16_t* su = img->as<uint16_t>();
      int lo_i = sg ? ss[0] : su[0], hi_i = lo_i;
      size_t i0 = 0;
#if defined(__SSE2__)
      if (n >= 8) {   // signed 16-bit min/max, eight at a time (unsigned samples are biased into signed order)
        const __m128i flip = _mm_set1_epi16(sg ? 0 : (short)0x8000);
        __m128i mn = _mm_xor_si128(_mm_loadu_si128(reinterpret_cast<const __m128i*>(ss)), flip), mx = mn;
        for (i0 = 8; i0 + 8 <= n; i0 += 8) {
          const __m128i v = _mm_xor_si128(_mm_loadu_si128(reinterpret_cast<const __m128i*>(ss + i0)), flip);
          mn = _mm_min_epi16(mn, v);
          mx = _mm_max_epi16(mx, v);
        }
        alignas(16) int16_t a[8], b[8];
        _mm_store_si128(reinterpret_cast<__m128i*>(a), mn);
        _mm_store_si128(reinterpret_cast<__m128i*>(b), mx);
        int lo_s = a[0], hi_s = b[0];
        for (int k = 1; k < 8; ++k) { lo_s = std::min<int>(lo_s, a[k]); hi_s = std::max<int>(hi_s, b[k]); }
        lo_i = sg ? lo_s : lo_s + 32768;
        hi_i = sg ? hi_s : hi_s + 32768;
      }
#endif
      if (sg) for (size_t i = i0; i < n; ++i) { lo_i = std::min<int>(lo_i, ss[i]); hi_i = std::max<int>(hi_i, ss[i]); }
      else for (size_t i = i0; i < n; ++i) { lo_i = std::min<int>(lo_i, su[i]); hi_i = std::max<int>(hi_i, su[i]); }
      const float lo = (float)lo_i, hi = (float)hi_i;
      const float inv = hi > lo ? 255.0f / (hi - lo) : 0.0f;
      std::vector<uint8_t> lut((size_t)(hi_i - lo_i) + 1);
      for (size_t k = 0; k < lut.size(); ++k) {
        const float g = ((float)(lo_i + (int)k) - lo) * inv;
        lut[k] = (uint8_t)(int)(std::min(255.0f, std::max(0.0f, g)) + 0.5f);
      }
      size_t i = 0;
      if (stride == 1) {
#if defined(__SSE2__)
        // eight pixels per step with the formula of the table itself (convert, subtract, scale, clamp, round half up,
        // truncate): the same IEEE operations per pixel, hence the same bytes as the table look-up
        const __m128 vlo = _mm_set1_ps(lo), vinv = _mm_set1_ps(inv), v255 = _mm_set1_ps(255.0f), v0 = _mm_setzero_ps(), vh = _mm_set1_ps(0.5f);
        const __m128i zero = _mm_setzero_si128();
        for (; i + 8 <= n; i += 8) {
          const __m128i s16 = _mm_loadu_si128(reinterpret_cast<const __m128i*>(ss + i));
          const __m128i ext = sg ? _mm_srai_epi16(s16, 15) : zero;
          __m128 g0 = _mm_mul_ps(_mm_sub_ps(_mm_cvtepi32_ps(_mm_unpacklo_epi16(s16, ext)), vlo), vinv);
          __m128 g1 = _mm_mul_ps(_mm_sub_ps(_mm_cvtepi32_ps(_mm_unpackhi_epi16(s16, ext)), vlo), vinv);
          g0 = _mm_add_ps(_mm_min_ps(_mm_max_ps(g0, v0), v255), vh);
          g1 = _mm_add_ps(_mm_min_ps(_mm_max_ps(g1, v0), v255), vh);
          const __m128i p16 = _mm_packs_epi32(_mm_cvttps_epi32(g0), _mm_cvttps_epi32(g1));
          _mm_storel_epi64(reinterpret_cast<__m128i*>(rgb + i), _mm_packus_epi16(p16, p16));
        }
#endif
        for (; i < n; ++i) rgb[i] = lut[(sg ? (int)ss[i] : (int)su[i]) - lo_i];
        return;
      }
      for (; i + 4 <= n; i += 4) {   // four pixels = twelve canvas bytes, written as three words
        uint32_t g[4];
        for (int k = 0; k < 4; ++k) g[k] = lut[(sg ? (int)ss[i + k] : (int)su[i + k]) - lo_i];
        const uint32_t w0 = g[0] * 0x010101u | g[1] << 24, w1 = g[1] * 0x0101u | g[2] * 0x01010000u, w2 = g[2] | g[3] * 0x01010100u;
        std::memcpy(rgb + 3 * i, &w0, 4); std::memcpy(rgb + 3 * i + 4, &w1, 4); std::memcpy(rgb + 3 * i + 8, &w2, 4);
      }
      for (; i < n; ++i) {
        const uint8_t g = lut[(sg ? (int)ss[i] : (int)su[i]) - lo_i];
        rgb[3 * i] = rgb[3 * i + 1] = rgb[3 * i + 2] = g;
      }
      return;
    }
    std::vector<float> v(n);  // one conversion pass, then plain float indexing
    switch (img->getDataType()) {
      case TYPE_FLOAT: std::memcpy(v.data(), img->as<float>(), n * 4); break;
      case TYPE_UINT8: { const uint8_t* s = img->as<uint8_t>(); for (size_t i = 0; i < n; ++i) v[i] = s[i]; break; }
      case TYPE_UINT16: { const uint16_t* s = img->as<uint16_t>(); for (size_t i = 0; i < n; ++i) v[i] = s[i]; break; }
      default: { const int16_t* s = img->as<int16_t>(); for (size_t i = 0; i < n; ++i) v[i] = s[i]; break; }
    }
    float lo = v[0], hi = v[0];  // window = full intensity range of the image, level = its centre
    for (size_t i = 0; i < n; ++i) { lo = std::min(lo, v[i]); hi = std::max(hi, v[i]); }
    const float inv = hi > lo ? 255.0f / (hi - lo) : 0.0f;
    auto grey = [&](float x) { const float g = (x - lo) * inv; return (uint8_t)(int)(std::min(255.0f, std::max(0.0f, g)) + 0.5f); };
    if (w == cw && h == ch) {  // canvas and image have the same size: the fit is the identity
      for (size_t i = 0; i < n; ++i) { const uint8_t g = grey(v[i]); for (int k = 0; k < stride; ++k) rgb[stride * i + k] = g; }
      return;
    }
    const Fit f = fit(w, h, cw, ch);
    const float is = 1.0f / f.scale;
    for (int cy = 0; cy < ch; ++cy) {
      const float sy = (cy + 0.5f - f.oy) * is - 0.5f;
      if (sy < -0.5f || sy > h - 0.5f) continue;
      const int y0 = std::max(0, std::min(h - 1, (int)std::floor(sy))), y1 = std::min(h - 1, y0 + 1);
      const float ay = std::min(1.0f, std::max(0.0f, sy - y0));
      for (int cx = 0; cx < cw; ++cx) {
        const float sx = (cx + 0.5f - f.ox) * is - 0.5f;
        if (sx < -0.5f || sx > w - 0.5f) continue;
        const int x0 = std::max(0, std::min(w - 1, (int)std::floor(sx))), x1 = std::min(w - 1, x0 + 1);
        const float ax = std::min(1.0f, std::max(0.0f, sx - x0));
        const float top = v[(size_t)y0 * w + x0] * (1 - ax) + v[(size_t)y0 * w + x1] * ax;
        const float bot = v[(size_t)y1 * w + x0] * (1 - ax) + v[(size_t)y1 * w + x1] * ax;
        const uint8_t g = grey(top * (1 - ay) + bot * ay);
        uint8_t* px = rgb + ((size_t)cy * cw + cx) * stride;
        for (int k = 0; k < stride; ++k) px[k] = g;
      }
    }
  }
 private:
  std::shared_ptr<ImageRenderer> me() { return std::static_pointer_cast<ImageRenderer>(shared_from_this()); }
};

class SegmentationRenderer : public Renderer {
 public:
  // SegmentationRenderer::create(labelColors, 0.6f, 1.0f, 2)  main_sequential.cpp:261
  static std::shared_ptr<SegmentationRenderer> create(LabelColors colors = {}, float opacity = 0.5f, float borderOpacity = 0.5f,
                                                      int borderRadius = 1) {
    return std::shared_ptr<SegmentationRenderer>(new SegmentationRenderer(std::move(colors), opacity, borderOpacity, borderRadius));
  }
  std::shared_ptr<SegmentationRenderer> connect(std::shared_ptr<ProcessObject> p) { connectParent(std::move(p)); return me(); }
  std::shared_ptr<SegmentationRenderer> addInputData(std::shared_ptr<Image> d) { connectData(std::move(d)); return me(); }
  void paint(uint8_t* rgb, int stride, int cw, int ch) override {
    auto img = inputImage();
    if (img->getDataType() != TYPE_UINT8) throw Exception("SegmentationRenderer needs a uint8 label image");
    const int w = img->getWidth(), h = img->getHeight();
    const Color* table[256] = {};
    for (auto& kv : colors_) if (kv.first >= 0 && kv.first < 256) table[kv.first] = &kv.second;
    if (w == cw && h == ch && std::max(0, radius_) <= 15 && paintBits(*img, table, rgb, stride)) return;
    const uint8_t* lab = img->as<uint8_t>();
    const Fit f = fit(w, h, cw, ch);
    // A pixel is a border pixel when some pixel within borderRadius (disc, outside the image = label 0) carries
    // another label.  Per labelled pixel, runr = how far its label extends unbroken to both sides in its row
    // (capped at the radius); the disc test is then one comparison per row of the disc.
    const int R = std::max(0, std::min(radius_, 127));
    std::vector<int> hw(2 * R + 1);
    for (int b = -R; b <= R; ++b) { int a = 0; while ((a + 1) * (a + 1) + b * b <= R * R) ++a; hw[b + R] = a; }
    std::vector<uint8_t> runr((size_t)w * h, 0);
    for (int y = 0; y < h; ++y) {
      const uint8_t* row = lab + (size_t)y * w;
      uint8_t* rr = runr.data() + (size_t)y * w;
      for (int x = 0; x < w; ++x) {
        if ((x & 7) == 0 && x + 8 <= w) {   // eight label-free pixels at once
          uint64_t eight;
          std::memcpy(&eight, row + x, 8);
          if (eight == 0) { x += 7; continue; }
        }
        const int l = row[x];
        if (l == 0) continue;
        int k = 0;
        while (k < R && x - k - 1 >= 0 && x + k + 1 < w && row[x - k - 1] == l && row[x + k + 1] == l) ++k;
        rr[x] = (uint8_t)k;
      }
    }
    auto is_border = [&](int x, int y, int l) {
      for (int b = -R; b <= R; ++b) {
        const int yy = y + b;
        if (yy < 0 || yy >= h) return true;
        const size_t i = (size_t)yy * w + x;
        if (lab[i] != l || runr[i] < hw[b + R]) return true;
      }
      return false;
    };
    std::vector<int> xs(cw);
    for (int cx = 0; cx < cw; ++cx) xs[cx] = (int)std::floor((cx + 0.5f - f.ox) / f.scale);
    for (int cy = 0; cy < ch; ++cy) {
      const int y = (int)std::floor((cy + 0.5f - f.oy) / f.scale);
      if (y < 0 || y >= h) continue;
      const uint8_t* row = lab + (size_t)y * w;
      const bool identity = cw == w && ch == h;   // canvas pixel = image pixel
      for (int cx = 0; cx < cw; ++cx) {
        if (identity && (cx & 7) == 0 && cx + 8 <= w) {
          uint64_t eight;
          std::memcpy(&eight, row + cx, 8);
          if (eight == 0) { cx += 7; continue; }
        }
        const int x = xs[cx];
        if (x < 0 || x >= w || row[x] == 0) continue;
        const int l = row[x];
        const Color* col = table[l];
        if (!col) continue;
        const float op = is_border(x, y, l) ? borderOpacity_ : opacity_;
        uint8_t* px = rgb + ((size_t)cy * cw + cx) * stride;
        const float c[3] = {col->r, col->g, col->b};
        for (int k = 0; k < stride; ++k) px[k] = (uint8_t)(int)(px[k] * (1 - op) + 255.0f * c[k] * op + 0.5f);
      }
    }
  }
  bool paintsGrey() const override {
    for (auto& kv : colors_) if (kv.second.r != kv.second.g || kv.second.g != kv.second.b) return false;
    return true;
  }
  // one channel of a label colour blended over a canvas byte with an opacity: the arithmetic of paint() (callers that
  // have the GPU render, nm03_outputs.export_seg_blocks, need the resulting greys)
  static uint8_t blend(uint8_t canvas, float colour, float opacity) {
    return (uint8_t)(int)(canvas * (1 - opacity) + 255.0f * colour * opacity + 0.5f);
  }
 private:
  // The driver's case — a {0,1} label image on a canvas of its own size — on bit planes: the labelled pixels as one bit
  // each (taken as they are from a bit-packed image, else packed here, giving up if any other label occurs), the
  // interior = erosion of that plane by the border disc with everything outside the image counting as label 0 (row by
  // row: AND of the row shifted by -a..a for the disc's half-width a at that row offset), and then the same blend as the
  // general path for every set bit, with the interior or the border opacity.  Same bytes, no per-pixel neighbourhood walk.
  bool paintBits(const Image& img, const Color* const* table, uint8_t* rgb, int stride) const {
    const int w = img.getWidth(), h = img.getHeight(), wpr = (w + 31) / 32;
    const Color* col = table[1];
    std::vector<uint32_t> packed;
    const uint32_t* L = img.labelBits();
    if (!L) {
      const uint8_t* lab = img.as<uint8_t>();
      packed.assign((size_t)h * wpr, 0u);
      for (int y = 0; y < h; ++y) {
        const uint8_t* row = lab + (size_t)y * w;
        uint32_t* out = packed.data() + (size_t)y * wpr;
        int x = 0;
#if defined(__SSE2__)
        const __m128i one = _mm_set1_epi8(1), zero = _mm_setzero_si128();
        for (; x + 16 <= w; x += 16) {
          const __m128i v = _mm_loadu_si128(reinterpret_cast<const __m128i*>(row + x));
          const int is1 = _mm_movemask_epi8(_mm_cmpeq_epi8(v, one)), is0 = _mm_movemask_epi8(_mm_cmpeq_epi8(v, zero));
          if ((is1 | is0) != 0xFFFF) return false;            // some other label: the general path handles it
          out[x >> 5] |= (uint32_t)is1 << (x & 31);
        }
#endif
        for (; x < w; ++x) {
          if (row[x] > 1) return false;
          out[x >> 5] |= (uint32_t)row[x] << (x & 31);
        }
      }
      L = packed.data();
    }
    if (!col) return true;                                     // label 1 has no colour: nothing is painted
    const int R = std::max(0, radius_);
    int hw[31];
    for (int b = -R; b <= R; ++b) { int a = 0; while ((a + 1) * (a + 1) + b * b <= R * R) ++a; hw[b + R] = a; }
    auto wordAt = [&](const uint32_t* row, int k) -> uint32_t { return (k < 0 || k >= wpr) ? 0u : row[k]; };
    // row eroded horizontally by half-width a (pixels outside the image are 0): word k of AND_{s=-a..a} (row >> s)
    auto erodeRow = [&](const uint32_t* row, int k, int a) -> uint32_t {
      uint32_t acc = row[k];
      const uint32_t lo = wordAt(row, k - 1), hi = wordAt(row, k + 1);
      for (int sft = 1; sft <= a && acc; ++sft)
        acc &= ((row[k] >> sft) | (hi << (32 - sft))) & ((row[k] << sft) | (lo >> (32 - sft)));
      return acc;
    };
    const float c[3] = {col->r, col->g, col->b};
    for (int y = 0; y < h; ++y) {
      const uint32_t* row = L + (size_t)y * wpr;
      for (int k = 0; k < wpr; ++k) {
        uint32_t m = row[k];
        if (!m) continue;
        uint32_t interior = m;
        for (int b = -R; b <= R && interior; ++b) {
          const int yy = y + b;
          if (yy < 0 || yy >= h) { interior = 0u; break; }
          interior &= erodeRow(L + (size_t)yy * wpr, k, hw[b + R]);
        }
        while (m) {
          const int bit = __builtin_ctz(m);
          m &= m - 1;
          const float op = (interior >> bit & 1u) ? opacity_ : borderOpacity_;
          uint8_t* px = rgb + ((size_t)y * w + 32 * k + bit) * stride;
          for (int q = 0; q < stride; ++q) px[q] = (uint8_t)(int)(px[q] * (1 - op) + 255.0f * c[q] * op + 0.5f);
        }
      }
    }
    return true;
  }
  SegmentationRenderer(LabelColors c, float o, float bo, int r) : colors_(std::move(c)), opacity_(o), borderOpacity_(bo), radius_(r) {}
  std::shared_ptr<SegmentationRenderer> me() { return std::static_pointer_cast<SegmentationRenderer>(shared_from_this()); }
  LabelColors colors_;
  float opacity_, borderOpacity_;
  int radius_;
};

class RenderToImage : public ProcessObject {
 public:
  static std::shared_ptr<RenderToImage> create(Color bg = Color::Black(), int width = 512, int height = 512) {
    return std::shared_ptr<RenderToImage>(new RenderToImage(bg, width, height));
  }
  std::shared_ptr<RenderToImage> connect(std::shared_ptr<Renderer> r) { renderers_.push_back(std::move(r)); output_.reset(); return me(); }
  void removeAllRenderers() { renderers_.clear(); output_.reset(); }
  void update() override { evaluate(); }
 protected:
  void evaluate() override {
    const uint8_t bg[3] = {(uint8_t)std::lround(bg_.r * 255), (uint8_t)std::lround(bg_.g * 255), (uint8_t)std::lround(bg_.b * 255)};
    bool grey = bg[0] == bg[1] && bg[1] == bg[2];
    for (auto& r : renderers_) grey = grey && r->paintsGrey();
    if (grey) {
      // every layer is grey (the reference's two exports are): one channel is painted, and the interleaved RGB bytes
      // are only produced if somebody reads them (the JPEG writer takes the single channel)
      std::vector<uint8_t> plane((size_t)w_ * h_, bg[0]);
      for (auto& r : renderers_) r->paint(plane.data(), 1, w_, h_);
      output_ = Image::greyCanvas(w_, h_, std::move(plane));
      return;
    }
    auto img = std::make_shared<Image>(w_, h_, TYPE_UINT8, 3);
    uint8_t* rgb = img->as<uint8_t>();
    for (size_t i = 0; i < (size_t)w_ * h_; ++i) { rgb[3 * i] = bg[0]; rgb[3 * i + 1] = bg[1]; rgb[3 * i + 2] = bg[2]; }
    for (auto& r : renderers_) r->paint(rgb, 3, w_, h_);
    output_ = img;
  }
 private:
  RenderToImage(Color bg, int w, int h) : ProcessObject(RENDER_TO_IMAGE), bg_(bg), w_(w), h_(h) {}
  std::shared_ptr<RenderToImage> me() { return std::static_pointer_cast<RenderToImage>(shared_from_this()); }
  Color bg_;
  int w_, h_;
  std::vector<std::shared_ptr<Renderer>> renderers_;
};

class ImageFileExporter : public Typed<ImageFileExporter> {
 public:
  // FAST's exporter hands a QImage to Qt's writer without a quality argument (QImage::save(filename)); Qt's JPEG
  // handler then uses its default, quality 75 [upstream recall, not verifiable offline].  The same default here;
  // setQuality() is an extension for callers that want something else.
  static constexpr int kDefaultJpegQuality = 75;
  void setQuality(int q) { quality_ = q < 1 ? 1 : (q > 100 ? 100 : q); }
  void update() override {
    auto img = inputImage();
    if (img->getDataType() != TYPE_UINT8) throw Exception("ImageFileExporter: only 8-bit images can be written as JPEG");
    const int ch = img->getNrOfChannels();
    try {
      if (const uint8_t* plane = img->greyPlane()) {   // a grey canvas of RenderToImage: one component, no conversion
        jpeg::save(file_, plane, img->getWidth(), img->getHeight(), 1, quality_);
        return;
      }
      if (ch == 3) {  // any other RGB image: store one component when all three agree
        const uint8_t* p = img->as<uint8_t>();
        const size_t n = (size_t)img->getWidth() * img->getHeight();
        bool grey = true;
        for (size_t i = 0; i < n && grey; ++i) grey = p[3 * i] == p[3 * i + 1] && p[3 * i] == p[3 * i + 2];
        if (grey) {
          std::vector<uint8_t> g(n);
          for (size_t i = 0; i < n; ++i) g[i] = p[3 * i];
          jpeg::save(file_, g.data(), img->getWidth(), img->getHeight(), 1, quality_);
          return;
        }
      }
      jpeg::save(file_, img->as<uint8_t>(), img->getWidth(), img->getHeight(), ch, quality_);
    } catch (const std::exception& e) {
      throw Exception(e.what());
    }
  }
 protected:
  friend class Typed<ImageFileExporter>;
  explicit ImageFileExporter(std::string file) : Typed(EXPORTER), file_(std::move(file)) {}
  std::string file_;
  int quality_ = kDefaultJpegQuality;
};

// The driver's second export in one step: SegmentationRenderer(label 1 = colour, opacity, borderOpacity, borderRadius) of
// a bit-packed {0,1} label image over a uniform grey background on a canvas of the image's own size, written as the
// baseline JPEG ImageFileExporter would write from that canvas — without the canvas.  A rendered mask is flat almost
// everywhere, and a flat block's only non-zero coefficient is known from its grey: blocks without a labelled pixel, and
// blocks that lie wholly in the interior, cost a few loads and six bits; only blocks the region's outline runs through
// are rendered, transformed and quantised.  Same bytes as the canvas path (tests/test_host_cpu.py).  w and h multiples
// of 8; bits as in nm03_outputs.dilated_bits.
inline std::vector<uint8_t> encodeSegmentationJpeg(const uint32_t* bits, int w, int h, float colour, float opacity, float borderOpacity,
                                                   int borderRadius, uint8_t background, int quality) {
  using namespace jpeg;
  using namespace jpeg::detail;
  if (w <= 0 || h <= 0 || (w & 7) || (h & 7)) throw Exception("encodeSegmentationJpeg: width and height must be multiples of 8");
  const int wpr = (w + 31) / 32, R = std::max(0, std::min(borderRadius, 15));
  const uint8_t grey_in = SegmentationRenderer::blend(background, colour, opacity);
  const uint8_t grey_bd = SegmentationRenderer::blend(background, colour, borderOpacity);
  // interior = the labelled pixels eroded by the border disc, everything outside the image unlabelled (paintBits)
  int hw[31];
  for (int b = -R; b <= R; ++b) { int a = 0; while ((a + 1) * (a + 1) + b * b <= R * R) ++a; hw[b + R] = a; }
  std::vector<uint32_t> rowe((size_t)(R + 1) * h * wpr);   // rowe[a][y][k]: row y eroded horizontally by half-width a
  for (int y = 0; y < h; ++y) {
    const uint32_t* row = bits + (size_t)y * wpr;
    for (int k = 0; k < wpr; ++k) {
      if (!row[k]) continue;                                   // (the planes start out zero)
      const uint32_t lo = k > 0 ? row[k - 1] : 0u, hi = k + 1 < wpr ? row[k + 1] : 0u;
      uint32_t acc = row[k];
      rowe[(size_t)y * wpr + k] = acc;
      for (int a = 1; a <= R; ++a) {
        acc &= ((row[k] >> a) | (hi << (32 - a))) & ((row[k] << a) | (lo >> (32 - a)));
        rowe[((size_t)a * h + y) * wpr + k] = acc;
      }
    }
  }
  std::vector<uint32_t> interior((size_t)h * wpr);
  for (int y = 0; y < h; ++y)
    for (int k = 0; k < wpr; ++k) {
      uint32_t acc = bits[(size_t)y * wpr + k];
      if (!acc) continue;
      for (int b = -R; b <= R && acc; ++b) {
        const int yy = y + b;
        acc = (yy < 0 || yy >= h) ? 0u : (acc & rowe[((size_t)hw[b + R] * h + yy) * wpr + k]);
      }
      interior[(size_t)y * wpr + k] = acc;
    }
  const Tables tb = make_tables(quality);
  const HuffSet& hf = huffman();
  std::vector<uint8_t> o;
  o.reserve((size_t)w * h / 16 + 1024);
  write_headers(o, w, h, 1, tb.q);
  auto flat_dc = [&](uint8_t g) {   // a flat block's DC term, as encode() computes it
    const float qv = ((float)g - 128.0f) * 64.0f * tb.rq[0][0];
    return (int16_t)(int)(qv + (qv >= 0.0f ? 0.5f : -0.5f));
  };
  const int16_t dc_bg = flat_dc(background), dc_in = flat_dc(grey_in);
  // a flat block whose DC term equals its predecessor's: the code of size 0, then the end-of-block code
  const int flat_same_len = hf.dc[0].len[0] + hf.ac[0].len[0x00];
  const uint32_t flat_same_bits = ((uint32_t)hf.dc[0].code[0] << hf.ac[0].len[0x00]) | hf.ac[0].code[0x00];
  const int bpr = w / 8;
  size_t pos = o.size();
  int pred = 0;
  BitWriter bw(nullptr);
  alignas(16) int16_t zz[64];
  std::memset(zz, 0, sizeof(zz));
  std::vector<uint32_t> any_w(wpr), all_w(wpr);
  static const auto spread = [] {   // bit i of the index -> byte i = 1
    std::vector<uint64_t> t(256);
    for (int b = 0; b < 256; ++b) { uint64_t v = 0; for (int i = 0; i < 8; ++i) v |= (uint64_t)((b >> i) & 1) << (8 * i); t[b] = v; }
    return t;
  }();
  for (int by = 0; by < h; by += 8) {
    if (o.size() - pos < (size_t)bpr * kWorstBlock + 64) o.resize(pos + (size_t)bpr * kWorstBlock * 2 + 64);
    bw.p = o.data() + pos;
    const uint8_t* mrow[8];
    const uint8_t* irow[8];
    for (int y = 0; y < 8; ++y) {   // byte bx of a row's words = the eight pixels of block column bx (little-endian words)
      mrow[y] = reinterpret_cast<const uint8_t*>(bits + (size_t)(by + y) * wpr);
      irow[y] = reinterpret_cast<const uint8_t*>(interior.data() + (size_t)(by + y) * wpr);
    }
    // the eight rows of the block row folded word by word: a byte of any_w is zero where a block has no labelled pixel, a
    // byte of all_w is 0xFF where every pixel of a block is an interior pixel (interior pixels are labelled)
    for (int k = 0; k < wpr; ++k) {
      uint32_t a = 0u, b = 0xFFFFFFFFu;
      for (int y = 0; y < 8; ++y) { a |= bits[(size_t)(by + y) * wpr + k]; b &= interior[(size_t)(by + y) * wpr + k]; }
      any_w[k] = a; all_w[k] = b;
    }
    const uint8_t* any_b = reinterpret_cast<const uint8_t*>(any_w.data());
    const uint8_t* all_b = reinterpret_cast<const uint8_t*>(all_w.data());
    for (int bx = 0; bx < bpr; ++bx) {
      if (any_b[bx] == 0 || all_b[bx] == 0xFF) {
        zz[0] = any_b[bx] == 0 ? dc_bg : dc_in;
        if (zz[0] == pred) bw.put(flat_same_bits, flat_same_len);   // DC difference 0, then end of block
        else put_block(bw, hf.dc[0], hf.ac[0], zz, 0, pred);
        continue;
      }
      alignas(8) uint8_t px[64];
      for (int y = 0; y < 8; ++y) {   // eight pixels at a time: the three pixel classes are disjoint, so the byte-wise products add without carries
        const uint8_t m = mrow[y][bx], in = irow[y][bx];
        const uint64_t row8 = spread[(uint8_t)~m] * background + spread[(uint8_t)(m & ~in)] * grey_bd + spread[in] * grey_in;
        std::memcpy(px + 8 * y, &row8, 8);
      }
      bool flat = true;
      for (int i = 1; i < 64 && flat; ++i) flat = px[i] == px[0];
      alignas(16) int16_t bz[64];
      uint64_t nz = 0;
      if (flat) { std::memset(bz, 0, sizeof(bz)); bz[0] = flat_dc(px[0]); }   // e.g. a block of border pixels only
      else block_coefficients(px, 8, tb.rq[0], bz, &nz);
      put_block(bw, hf.dc[0], hf.ac[0], bz, nz, pred);
    }
    pos = (size_t)(bw.p - o.data());
  }
  if (o.size() - pos < 32) o.resize(pos + 32);
  bw.p = o.data() + pos;
  bw.flush();
  pos = (size_t)(bw.p - o.data());
  o.resize(pos);
  o.push_back(0xFF); o.push_back(0xD9);  // EOI
  return o;
}

// The 5-pane GUI of test_pipeline.cpp:148-158 needs a window system; it is left out (no graphics
// stack on this box).  run() returns immediately so the export stage that follows still executes.
class MultiViewWindow {
 public:
  static std::shared_ptr<MultiViewWindow> create(int, Color = Color::Black(), int = 0, int = 0, bool = false) {
    return std::make_shared<MultiViewWindow>();
  }
  void addRenderer(int, std::shared_ptr<Renderer>) {}
  void setTitle(const std::string&) {}
  void run() {}
};

}  // namespace nm03
