// nm03_driver.cpp — B200-native batch driver with the shape of the reference's executables.
//
//   --mode parallel    img_processing_parallel   (src/parallel/main_parallel.cpp): patients sharded over
//                      the GPUs by contiguous ranges, one feeder per GPU stream slot, whole-patient
//                      batches through nm03_process_batch, JPEG export on host worker threads
//   --mode sequential  img_processing_sequential (src/sequential/main_sequential.cpp): one slice at a time
//   --test-pipeline F  test_pipeline            (src/test/test_pipeline.cpp): one slice, erosion branch,
//                      five JPEGs (the GUI window is left out: no graphics stack)
//   --make-dataset D   writes a synthetic cohort  D/PGBM-%03d/<series>/1-%02d.dcm  
//
// Kept from the reference: PGBM-* discovery (main_sequential.cpp:93-119), first series directory
// (:138-139), numeric file ordering with fallback 1000 (:18-30,152-154), clean per-patient output
// directory (:32-47), <stem>_original.jpg / <stem>_processed.jpg (:55-73), catch-and-continue with
// per-patient success counts (:286-299).  Not kept: system("mkdir/rm") (std::filesystem instead).
#include <nm03.h>

#include <malloc.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <condition_variable>
#include <deque>
#include <filesystem>
#include <functional>
#include <iostream>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "nm03.hpp"
#include "phantom.hpp"

namespace fs = std::filesystem;
using Clock = std::chrono::steady_clock;

namespace {

struct Options {
  std::string mode = "parallel", data, out = "../out-parallel", test_file, make_dataset;
  std::vector<int> gpus;
  int batch = 25, streams = 0, host_threads = 0;   // feeders per GPU, 0 = choose (see runBatch)
  bool do_export = true, dump_masks = false, lowcontrast = false;
  // --gpu-seg: the RESULT's rendering, transform and quantisation on the GPU as well.  Off by default: its blocks double the
  // bytes a batch brings back, and on one GPU + 16 cores that costs the feeders more than the host saves (20 100 vs 21 800
  // slices/s with 3 feeders; all on the host: 15 300)
  bool gpu_seg = false;
  bool cpu_dct = false;      // --cpu-dct: render, transform and quantise the original on the host (default: the GPU does, next to the pipeline)
  bool byte_masks = false;   // --byte-masks: fetch the uint8 mask from the GPU instead of the bit-packed one (8x the bytes over the link)
  int patients = 20, slices = 24, size = 512;
};

// main_sequential.cpp:18-30
int extractFileNumber(const std::string& filename) {
  const size_t dash = filename.find_last_of('-'), dot = filename.find(".dcm");
  if (dash != std::string::npos && dot != std::string::npos && dot > dash) {
    try { return std::stoi(filename.substr(dash + 1, dot - dash - 1)); } catch (...) { return 1000; }
  }
  return 1000;
}

std::vector<std::string> findAllPatientDirectories(const std::string& base) {
  std::vector<std::string> dirs;
  for (const auto& e : fs::directory_iterator(base))
    if (e.is_directory() && e.path().filename().string().rfind("PGBM-", 0) == 0) dirs.push_back(e.path().filename().string());
  std::sort(dirs.begin(), dirs.end());
  return dirs;
}

std::vector<std::string> loadDICOMFilesForPatient(const std::string& base, const std::string& patient) {
  std::vector<std::string> series;
  for (const auto& e : fs::directory_iterator(fs::path(base) / patient))
    if (e.is_directory()) series.push_back(e.path().string());
  if (series.empty()) throw std::runtime_error("No series directories found for patient: " + patient);
  std::sort(series.begin(), series.end());  // the reference takes whatever directory_iterator yields first
  std::vector<std::pair<std::string, int>> files;
  for (const auto& e : fs::directory_iterator(series[0]))
    if (e.path().extension() == ".dcm") files.push_back({e.path().string(), extractFileNumber(e.path().filename().string())});
  std::stable_sort(files.begin(), files.end(), [](const auto& a, const auto& b) { return a.second < b.second; });
  std::vector<std::string> out;
  for (auto& f : files) out.push_back(f.first);
  return out;
}

void setupOutputDirectory(const fs::path& dir) {
  fs::create_directories(dir);
  for (const auto& e : fs::directory_iterator(dir)) fs::remove_all(e.path());
}

// persistent host worker pool shared by all feeders (parsing, staging copies, JPEG export)
class Pool {
 public:
  explicit Pool(int threads) {
    for (int t = 0; t < threads; ++t)
      workers_.emplace_back([this] {
        for (;;) {
          std::function<void()> job;
          {
            std::unique_lock<std::mutex> lk(mu_);
            cv_.wait(lk, [this] { return stop_ || !jobs_.empty(); });
            if (stop_ && jobs_.empty()) return;
            job = std::move(jobs_.front());
            jobs_.pop_front();
          }
          job();
        }
      });
  }
  ~Pool() {
    { std::lock_guard<std::mutex> lk(mu_); stop_ = true; }
    cv_.notify_all();
    for (auto& t : workers_) t.join();
  }
  // run f(0..n-1) on the pool (the caller works too) and wait
  template <class F>
  void parallel_for(int n, F&& f) {
    if (n <= 0) return;
    struct Shared { std::atomic<int> next{0}, done{0}; std::mutex m; std::condition_variable cv; };
    auto sh = std::make_shared<Shared>();
    auto body = [sh, n, &f] {
      for (int i; (i = sh->next.fetch_add(1)) < n;) {
        f(i);
        if (sh->done.fetch_add(1) + 1 == n) { std::lock_guard<std::mutex> lk(sh->m); sh->cv.notify_all(); }
      }
    };
    const int helpers = std::min<int>((int)workers_.size(), n - 1);
    {
      std::lock_guard<std::mutex> lk(mu_);
      for (int t = 0; t < helpers; ++t) jobs_.push_front(body);   // ahead of queued export work: a feeder is waiting for this
    }
    cv_.notify_all();
    body();
    std::unique_lock<std::mutex> lk(sh->m);
    sh->cv.wait(lk, [&] { return sh->done.load() >= n; });
  }
  // run f(0..n-1) on the pool without waiting; on_done() runs on the worker that finishes the last index.
  // Queued behind everything else: this is the export stage, which nobody waits for until the run ends.
  void submit_batch(int n, std::function<void(int)> f, std::function<void()> on_done) {
    { std::lock_guard<std::mutex> lk(mu_); outstanding_ += 1; }
    if (n <= 0 || workers_.empty()) { for (int i = 0; i < n; ++i) f(i); on_done(); return; }
    struct Shared { std::atomic<int> next{0}, done{0}; int n; std::function<void(int)> f; std::function<void()> on_done; };
    auto sh = std::make_shared<Shared>();
    sh->n = n; sh->f = std::move(f); sh->on_done = std::move(on_done);
    auto body = [sh] {
      for (int i; (i = sh->next.fetch_add(1)) < sh->n;) {
        sh->f(i);
        if (sh->done.fetch_add(1) + 1 == sh->n) sh->on_done();
      }
    };
    {
      std::lock_guard<std::mutex> lk(mu_);
      const int copies = std::min<int>((int)workers_.size(), n);
      for (int t = 0; t < copies; ++t) jobs_.push_back(body);
    }
    cv_.notify_all();
  }
  void batch_finished() {
    { std::lock_guard<std::mutex> lk(mu_); outstanding_ -= 1; }
    idle_cv_.notify_all();
  }
  void wait_batches() {
    std::unique_lock<std::mutex> lk(mu_);
    idle_cv_.wait(lk, [this] { return outstanding_ == 0; });
  }
 private:
  int outstanding_ = 0;
  std::condition_variable idle_cv_;
  std::vector<std::thread> workers_;
  std::deque<std::function<void()>> jobs_;
  std::mutex mu_;
  std::condition_variable cv_;
  bool stop_ = false;
};

Pool* g_pool = nullptr;

template <class F>
void parallel_for(int n, int threads, F&& f) {
  if (threads <= 1 || n <= 1 || !g_pool) { for (int i = 0; i < n; ++i) f(i); return; }
  g_pool->parallel_for(n, f);
}

struct Pinned {
  void* p = nullptr;
  size_t bytes = 0;
  void reserve(size_t n) {
    if (n <= bytes) return;
    if (p) nm03_host_free(p);
    if (nm03_host_alloc(&p, n) != NM03_OK) throw std::runtime_error(nm03_last_error());
    bytes = n;
  }
  ~Pinned() { if (p) nm03_host_free(p); }
};

// bit-packed mask row-words -> uint8 {0,1} pixels (nm03_outputs.dilated_bits; bit x%32 of word x/32 = pixel x)
void unpackMask(const uint32_t* bits, int w, int h, uint8_t* out) {
  static const auto table = [] {
    std::vector<uint64_t> t(256);
    for (int b = 0; b < 256; ++b) { uint64_t v = 0; for (int i = 0; i < 8; ++i) v |= (uint64_t)((b >> i) & 1) << (8 * i); t[b] = v; }
    return t;
  }();
  const int wpr = (w + 31) / 32;
  for (int y = 0; y < h; ++y) {
    const uint32_t* row = bits + (size_t)y * wpr;
    uint8_t* o = out + (size_t)y * w;
    int x = 0;
    for (int k = 0; k < wpr; ++k) {
      const uint32_t word = row[k];
      for (int byte = 0; byte < 4 && x < w; ++byte, x += 8) {
        const uint64_t v = table[(word >> (8 * byte)) & 0xffu];
        if (x + 8 <= w) std::memcpy(o + x, &v, 8);
        else std::memcpy(o + x, &v, (size_t)(w - x));
      }
    }
  }
}

// Pinned staging of one batch (input slices, result masks).  A feeder owns a small ring of them: a set is busy from
// the moment its batch is read until the export jobs that render from it have finished.
struct BufSet {
  Pinned in, out;
  bool busy = false;
};
struct BufRing {
  std::vector<std::unique_ptr<BufSet>> sets;
  std::mutex mu;
  std::condition_variable cv;
  BufSet* acquire() {
    std::unique_lock<std::mutex> lk(mu);
    for (;;) {
      for (auto& s : sets) if (!s->busy) { s->busy = true; return s.get(); }
      cv.wait(lk);
    }
  }
  void release(BufSet* s) {
    { std::lock_guard<std::mutex> lk(mu); s->busy = false; }
    cv.notify_one();
  }
};

// SegmentationRenderer::create(labelColors, 0.6f, 1.0f, 2) with label 1 = white over black, main_sequential.cpp:255-261
constexpr float kSegOpacity = 0.6f, kSegBorderOpacity = 1.0f;
constexpr int kSegBorderRadius = 2;
const nm03_export_style kSegStyle = {0, nm03::SegmentationRenderer::blend(0, 1.0f, kSegOpacity),
                                     nm03::SegmentationRenderer::blend(0, 1.0f, kSegBorderOpacity), kSegBorderRadius};

// mask: uint8 labels, or (bits = true) the bit-packed words of nm03_outputs.dilated_bits
// blocks / nz: the original's quantised 8x8 blocks from the GPU (nm03_outputs.export_blocks), or null: rendered and
// transformed here
void exportSlice(const fs::path& dir, const std::string& file, const int16_t* raw, bool is_unsigned, const void* mask, bool bits, int w, int h,
                 const int16_t* blocks = nullptr, const uint64_t* nz = nullptr, const int16_t* seg_blocks = nullptr,
                 const uint64_t* seg_nz = nullptr) {
  using namespace nm03;
  const std::string stem = fs::path(file).stem().string();
  // views of the pinned staging buffers (they outlive this call): no copies, and the segmentation renderer works on the
  // bit-packed words as they come from the GPU
  auto orig = Image::view(raw, w, h, is_unsigned ? TYPE_UINT16 : TYPE_INT16);
  auto seg = bits ? Image::viewBits((const uint32_t*)mask, w, h) : Image::view(mask, w, h, TYPE_UINT8);
  LabelColors colors;
  colors[1] = Color::White();
  auto rti = RenderToImage::create(Color::Black(), 512, 512);  // main_sequential.cpp:258
  // the canvas is the slice's own size, so the GPU's blocks are the rendering's: only the entropy coding is left
  auto write_blocks = [&](const int16_t* zz, const uint64_t* map, const char* suffix) {
    const std::vector<uint8_t> b = jpeg::encode_blocks(zz, map, w, h, ImageFileExporter::kDefaultJpegQuality);
    const std::string path = (dir / (stem + suffix)).string();
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) throw Exception("cannot write " + path);
    const size_t nw = std::fwrite(b.data(), 1, b.size(), f);
    std::fclose(f);
    if (nw != b.size()) throw Exception("short write " + path);
  };
  const bool canvas_is_slice = w == 512 && h == 512;
  if (blocks && canvas_is_slice) {
    write_blocks(blocks, nz, "_original.jpg");
  } else {
    rti->connect(ImageRenderer::create()->addInputData(orig));
    rti->update();
    ImageFileExporter::create((dir / (stem + "_original.jpg")).string())->connect(rti->getOutputData<Image>(0))->update();
  }
  if (seg_blocks && canvas_is_slice) { write_blocks(seg_blocks, seg_nz, "_processed.jpg"); return; }
  if (bits && canvas_is_slice) {
    // the result's rendering and its JPEG in one step from the bit-packed mask: no canvas, flat blocks short-cut
    const std::vector<uint8_t> b = encodeSegmentationJpeg((const uint32_t*)mask, w, h, 1.0f, kSegOpacity, kSegBorderOpacity, kSegBorderRadius,
                                                          0, ImageFileExporter::kDefaultJpegQuality);
    const std::string path = (dir / (stem + "_processed.jpg")).string();
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) throw Exception("cannot write " + path);
    const size_t nw = std::fwrite(b.data(), 1, b.size(), f);
    std::fclose(f);
    if (nw != b.size()) throw Exception("short write " + path);
    return;
  }
  rti->removeAllRenderers();
  rti->connect(SegmentationRenderer::create(colors, kSegOpacity, kSegBorderOpacity, kSegBorderRadius)->addInputData(seg));
  rti->update();
  ImageFileExporter::create((dir / (stem + "_processed.jpg")).string())->connect(rti->getOutputData<Image>(0))->update();
}

struct Totals {
  std::atomic<long> ok{0}, total{0}, patients_ok{0};
  std::atomic<long> ns_list{0}, ns_parse{0}, ns_read{0}, ns_gpu{0}, ns_export{0}, ns_submit{0};  // feeder wall time per phase (NM03_DRIVER_TIMING=1 prints it)
  std::mutex log;
};

struct PhaseTimer {   // adds the time since construction (or the last lap) to a counter
  Clock::time_point t = Clock::now();
  void lap(std::atomic<long>& acc) {
    const auto n = Clock::now();
    acc += std::chrono::duration_cast<std::chrono::nanoseconds>(n - t).count();
    t = n;
  }
};

// one patient: parse on host threads into pinned memory, one GPU call per group of equal-size slices, export
void processPatient(nm03_ctx* ctx, int gpu, const Options& opt, const std::string& patient, BufRing& ring, Totals& tot) {
  PhaseTimer timer;
  const fs::path outdir = fs::path(opt.out) / patient;
  if (opt.do_export || opt.dump_masks) setupOutputDirectory(outdir);
  const std::vector<std::string> files = loadDICOMFilesForPatient(opt.data, patient);
  const int n = (int)files.size();
  tot.total += n;
  timer.lap(tot.ns_list);
  // the header of every file is parsed from its first bytes; the pixel payload is then read straight into the
  // pinned staging buffer (one copy)
  std::vector<nm03::dicom::FileReader> readers(n);
  std::vector<nm03::dicom::View> slices(n);
  std::vector<int> state(n, 0);  // 0 = parsed, <0 = failed
  parallel_for(n, opt.host_threads, [&](int i) {
    try {
      readers[i] = nm03::dicom::FileReader(files[i]);
      slices[i] = readers[i].view();
    } catch (const std::exception& e) {
      state[i] = -1;
      std::lock_guard<std::mutex> lk(tot.log);
      std::cerr << "Error processing file " << files[i] << ":\nDetailed error: " << e.what() << std::endl;
    }
  });
  timer.lap(tot.ns_parse);
  nm03_params params;
  nm03_default_params(&params);
  int ok = 0;
  std::vector<char> done(n, 0);
  for (int first = 0; first < n; ++first) {
    if (done[first] || state[first] < 0) continue;
    // group: same size and signedness, in file order, at most `batch` slices (main_parallel.cpp:330-333)
    std::vector<int> idx;
    for (int j = first; j < n && (int)idx.size() < opt.batch; ++j)
      if (!done[j] && state[j] == 0 && slices[j].cols == slices[first].cols && slices[j].rows == slices[first].rows &&
          slices[j].pixel_representation == slices[first].pixel_representation &&
          slices[j].rescale_slope == slices[first].rescale_slope && slices[j].rescale_intercept == slices[first].rescale_intercept) { idx.push_back(j); done[j] = 1; }
    const int w = slices[first].cols, h = slices[first].rows, m = (int)idx.size();
    const size_t npx = (size_t)w * h;
    BufSet* bs = ring.acquire();          // waits while all sets of this feeder are still being exported from
    struct Release { BufRing& r; BufSet* b; ~Release() { if (b) r.release(b); } } rel{ring, bs};
    Pinned& in = bs->in;
    Pinned& out = bs->out;
    in.reserve(m * npx * 2);
    std::atomic<int> read_failed{0};
    parallel_for(m, opt.host_threads, [&](int j) {
      try {
        readers[idx[j]].read_at((uint8_t*)in.p + j * npx * 2, npx * 2, slices[idx[j]].pixel_offset);
        // modality LUT (identity for almost every MR file: then nothing is touched)
        nm03::dicom::rescale_in_place((uint16_t*)((uint8_t*)in.p + j * npx * 2), npx, slices[idx[j]].pixel_representation,
                                      slices[idx[j]].rescale_slope, slices[idx[j]].rescale_intercept);
      } catch (const std::exception& e) {
        read_failed = 1;
        std::lock_guard<std::mutex> lk(tot.log);
        std::cerr << "Error processing file " << files[idx[j]] << ":\nDetailed error: " << e.what() << std::endl;
      }
    });
    if (read_failed) continue;   // the group is skipped, the patient goes on (main_parallel.cpp:163-167)
    timer.lap(tot.ns_read);
    params.input_unsigned = nm03::dicom::rescaled_representation(slices[first].pixel_representation, slices[first].rescale_slope,
                                                                 slices[first].rescale_intercept) == 0;
    // the mask comes back bit-packed (1/8 of the bytes over the host link) and is unpacked by the worker that needs it
    const size_t wpr = (size_t)(w + 31) / 32, mask_bytes = opt.byte_masks ? npx : h * wpr * 4;
    // with export on 512x512 slices (the canvas size of the two renders) the GPU also hands back the original's
    // quantised JPEG blocks: 128 bytes of coefficients + an 8-byte map per 8x8 block, behind the masks
    const bool gpu_blocks = opt.do_export && !opt.cpu_dct && w == 512 && h == 512;
    const size_t nblk = (size_t)(w / 8) * (h / 8), masks_end = (m * mask_bytes + 63) & ~(size_t)63;
    const bool gpu_seg_blocks = gpu_blocks && !opt.byte_masks && opt.gpu_seg;   // the result's rendering too (it is made from the bit-packed mask)
    out.reserve(masks_end + (gpu_blocks ? m * nblk * 136 : 0) + (gpu_seg_blocks ? m * nblk * 136 : 0));
    nm03_outputs o = {};
    if (opt.byte_masks) o.dilated = (uint8_t*)out.p;
    else o.dilated_bits = (uint32_t*)out.p;
    static const nm03::jpeg::Tables tables = nm03::jpeg::make_tables(nm03::ImageFileExporter::kDefaultJpegQuality);
    if (gpu_blocks) {
      o.export_blocks = (int16_t*)((uint8_t*)out.p + masks_end);
      o.export_nz = (uint64_t*)((uint8_t*)out.p + masks_end + m * nblk * 128);
      o.export_rq = tables.rq[0];
    }
    if (gpu_seg_blocks) {
      o.export_seg_blocks = (int16_t*)((uint8_t*)out.p + masks_end + m * nblk * 136);
      o.export_seg_nz = (uint64_t*)((uint8_t*)out.p + masks_end + m * nblk * 136 + m * nblk * 128);
      o.export_seg_style = &kSegStyle;
    }
    std::vector<int> status(m, 0);
    const int rc = nm03_process_batch(ctx, gpu, in.p, m, w, h, &params, nullptr, 0, &o, status.data());
    timer.lap(tot.ns_gpu);
    if (rc != NM03_OK) {
      std::lock_guard<std::mutex> lk(tot.log);
      for (int j = 0; j < m; ++j)
        std::cerr << "Error processing file " << files[idx[j]] << ":\nDetailed error: " << nm03_last_error() << std::endl;
      continue;  // a failed group is skipped, the patient goes on (main_parallel.cpp:163-167)
    }
    ok += m;
    auto mask_of = [&](int j, std::vector<uint8_t>& scratch) -> const uint8_t* {
      if (opt.byte_masks) return (const uint8_t*)out.p + j * npx;
      scratch.resize(npx);
      unpackMask((const uint32_t*)((const uint8_t*)out.p + j * mask_bytes), w, h, scratch.data());
      return scratch.data();
    };
    if (opt.dump_masks)
      for (int j = 0; j < m; ++j) {
        std::vector<uint8_t> scratch;
        const uint8_t* mk = mask_of(j, scratch);
        FILE* f = std::fopen((outdir / (fs::path(files[idx[j]]).stem().string() + ".mask")).string().c_str(), "wb");
        if (f) { std::fwrite(mk, 1, npx, f); std::fclose(f); }
      }
    if (opt.do_export && g_pool && opt.host_threads > 1) {
      // The export of this batch is queued on the pool and the feeder moves on to its next patient; the staging set goes
      // back to the ring when the last slice has been written (the reference exports serially after each batch,
      // main_parallel.cpp:346 — here that work keeps every host core busy while the GPUs go on).
      std::vector<std::string> names;
      for (int j = 0; j < m; ++j) names.push_back(files[idx[j]]);
      const void* in_p = in.p;
      const void* out_p = out.p;
      const bool is_unsigned = params.input_unsigned != 0, byte_masks = opt.byte_masks;
      const int16_t* blk_p = o.export_blocks;
      const uint64_t* nz_p = o.export_nz;
      const int16_t* sblk_p = o.export_seg_blocks;
      const uint64_t* snz_p = o.export_seg_nz;
      Totals* totp = &tot;
      BufRing* ringp = &ring;
      Pool* pool = g_pool;
      rel.b = nullptr;                    // ownership of the set passes to the export jobs
      pool->submit_batch(m, [=](int j) {
        const auto t0 = Clock::now();
        try {
          exportSlice(outdir, names[j], (const int16_t*)((const uint8_t*)in_p + j * npx * 2), is_unsigned,
                      (const uint8_t*)out_p + j * mask_bytes, !byte_masks, w, h, blk_p ? blk_p + j * nblk * 64 : nullptr,
                      nz_p ? nz_p + j * nblk : nullptr, sblk_p ? sblk_p + j * nblk * 64 : nullptr, snz_p ? snz_p + j * nblk : nullptr);
        } catch (const std::exception& e) {
          std::lock_guard<std::mutex> lk(totp->log);
          std::cerr << "Error in export stage: " << e.what() << std::endl;
        }
        totp->ns_export += std::chrono::duration_cast<std::chrono::nanoseconds>(Clock::now() - t0).count();
      }, [=] { ringp->release(bs); pool->batch_finished(); });
      timer.lap(tot.ns_submit);
    } else if (opt.do_export) {
      for (int j = 0; j < m; ++j) {
        try {
          exportSlice(outdir, files[idx[j]], (const int16_t*)((uint8_t*)in.p + j * npx * 2), params.input_unsigned,
                      (const uint8_t*)out.p + j * mask_bytes, !opt.byte_masks, w, h,
                      o.export_blocks ? o.export_blocks + j * nblk * 64 : nullptr, o.export_nz ? o.export_nz + j * nblk : nullptr,
                      o.export_seg_blocks ? o.export_seg_blocks + j * nblk * 64 : nullptr,
                      o.export_seg_nz ? o.export_seg_nz + j * nblk : nullptr);
        } catch (const std::exception& e) {
          std::lock_guard<std::mutex> lk(tot.log);
          std::cerr << "Error in export stage: " << e.what() << std::endl;
        }
      }
      timer.lap(tot.ns_export);
    }
  }
  tot.ok += ok;
  tot.patients_ok += 1;
  std::lock_guard<std::mutex> lk(tot.log);
  std::cout << "Patient " << patient << " completed. Successfully processed " << ok << "/" << n << " images." << std::endl;
}

int runBatch(Options opt) {
  const auto t_start = Clock::now();
  const std::vector<std::string> patients = findAllPatientDirectories(opt.data);
  std::cout << "Found " << patients.size() << " patient directories." << std::endl;
  if (patients.empty()) return 0;
  fs::create_directories(opt.out);
  nm03_ctx* ctx = nullptr;
  std::vector<int> gpus = opt.gpus;
  if (gpus.empty() && opt.mode != "sequential")   // default: every visible GPU, patients sharded across them
    for (int g = 0; g < nm03_device_count(); ++g) gpus.push_back(g);
  if (nm03_init(gpus.empty() ? nullptr : gpus.data(), (int)gpus.size(), &ctx) != NM03_OK) {
    std::cerr << "Fatal error: " << nm03_last_error() << std::endl;
    return 1;
  }
  Pool pool(std::max(0, opt.host_threads - 1));
  g_pool = &pool;
  const int G = nm03_gpu_count(ctx), P = (int)patients.size();
  const bool sequential = opt.mode == "sequential";
  Totals tot;
  // Start-up, outside the timed region: the pinned staging ring (one in/out pair per feeder, sized for a batch of
  // 512x512 slices and grown later if a patient needs more) and one batch of blank slices per GPU, which loads the
  // kernels and sizes the library's device scratch.  Reported as startup_seconds.
  // Feeders per GPU.  Without export the feeders' blocking GPU calls should overlap each other: three.  With export the
  // feeders share the host with the export jobs and about three to eight of them in total are best (r2t: 1 GPU + 16 cores
  // 10.7 k / 14.5 k / 14.7 k / 13.2 k slices/s with 1 / 2 / 3 / 4 feeders; 8 GPUs + 32 cores 22.1 k / 16.0 k / 15.5 k with
  // 1 / 2 / 3 per GPU): three on one GPU, two each on two, one each beyond.
  if (opt.streams <= 0) opt.streams = !opt.do_export ? 3 : (G >= 3 ? 1 : (G == 2 ? 2 : 3));
  const int n_feeders_per_gpu = sequential ? 1 : std::max(1, opt.streams);
  const int n_gpus_used = sequential ? 1 : G;
  // three staging sets per feeder: one being read / on the GPU, two whose export jobs are still running
  std::vector<std::unique_ptr<BufRing>> rings;
  for (int i = 0; i < n_gpus_used * n_feeders_per_gpu; ++i) {
    rings.emplace_back(new BufRing());
    for (int b = 0; b < (opt.do_export && !sequential ? 3 : 1); ++b) {
      rings.back()->sets.emplace_back(new BufSet());
      rings.back()->sets.back()->in.reserve((size_t)opt.batch * 512 * 512 * 2);
      // masks (at most a byte per pixel) and, with export, the original's JPEG blocks from the GPU (136 bytes per 8x8 block)
      rings.back()->sets.back()->out.reserve((size_t)opt.batch * 512 * 512 + 64 +
                                             (opt.do_export && !opt.cpu_dct ? (size_t)opt.batch * 4096 * 136 * (opt.gpu_seg ? 2 : 1) : 0));
    }
  }
  for (int g = 0; g < n_gpus_used; ++g) {
    Pinned& in = rings[g * n_feeders_per_gpu]->sets[0]->in;
    Pinned& out = rings[g * n_feeders_per_gpu]->sets[0]->out;
    const int m = sequential ? 1 : opt.batch;
    std::memset(in.p, 0, (size_t)m * 512 * 512 * 2);
    nm03_params params;
    nm03_default_params(&params);
    nm03_outputs o = {};
    if (opt.byte_masks || !opt.do_export) o.dilated = (uint8_t*)out.p;
    else o.dilated_bits = (uint32_t*)out.p;
    static const nm03::jpeg::Tables warm_tables = nm03::jpeg::make_tables(nm03::ImageFileExporter::kDefaultJpegQuality);
    if (opt.do_export && !opt.cpu_dct) {   // as the feeders will ask for it, so that the staging is sized and the kernels are loaded
      o.export_blocks = (int16_t*)((uint8_t*)out.p + (size_t)m * 512 * 512 + 64);
      o.export_nz = (uint64_t*)((uint8_t*)out.p + (size_t)m * 512 * 512 + 64 + (size_t)m * 4096 * 128);
      o.export_rq = warm_tables.rq[0];
      if (!opt.byte_masks && opt.gpu_seg) {
        o.export_seg_blocks = (int16_t*)((uint8_t*)out.p + (size_t)m * 512 * 512 + 64 + (size_t)m * 4096 * 136);
        o.export_seg_nz = (uint64_t*)((uint8_t*)out.p + (size_t)m * 512 * 512 + 64 + (size_t)m * 4096 * 136 + (size_t)m * 4096 * 128);
        o.export_seg_style = &kSegStyle;
      }
    }
    if (nm03_process_batch(ctx, g, in.p, m, 512, 512, &params, nullptr, 0, &o, nullptr) != NM03_OK) {
      std::cerr << "Fatal error: " << nm03_last_error() << std::endl;
      nm03_shutdown(ctx);
      return 1;
    }
  }
  const auto t0 = Clock::now();
  const double startup = std::chrono::duration<double>(t0 - t_start).count();
  std::vector<std::thread> feeders;
  for (int g = 0; g < (sequential ? 1 : G); ++g) {
    // contiguous patient range of GPU g
    const int lo = sequential ? 0 : (int)((long)g * P / G), hi = sequential ? P : (int)((long)(g + 1) * P / G);
    // Feeders per GPU: each takes whole patients and spreads a patient's slices over the shared pool.  Measured on 120
    // patients x 24 slices in /dev/shm, 16 cores: 3 feeders 21-29k slices/s, 8 feeders 18k, 16 pool-less feeders 8-11k
    // (every feeder then pins its own staging buffers and reads its 24 files one after the other).
    const int S = sequential ? 1 : std::max(1, opt.streams);
    const int inner_threads = opt.host_threads;
    auto next = std::make_shared<std::atomic<int>>(lo);
    for (int s = 0; s < S; ++s)
      feeders.emplace_back([&, g, s, hi, next, inner_threads] {
        BufRing& ring = *rings[g * n_feeders_per_gpu + s];
        for (int i; (i = next->fetch_add(1)) < hi;) {
          try {
            if (sequential) {
              Options one = opt;
              one.batch = 1;  // one slice per GPU call, as the reference's serial loop does
              one.host_threads = 1;
              processPatient(ctx, g, one, patients[i], ring, tot);
            } else {
              Options mine = opt;
              mine.host_threads = inner_threads;
              processPatient(ctx, g, mine, patients[i], ring, tot);
            }
          } catch (const std::exception& e) {
            std::lock_guard<std::mutex> lk(tot.log);
            std::cerr << "Error processing patient " << patients[i] << ": " << e.what() << std::endl;
          }
        }
      });
  }
  for (auto& t : feeders) t.join();
  pool.wait_batches();                 // the export jobs of the last batches
  g_pool = nullptr;
  const double sec = std::chrono::duration<double>(Clock::now() - t0).count();
  std::cout << "\n=== All Processing Completed ===\n\nSuccessfully processed " << tot.patients_ok << "/" << P << " patients." << std::endl;
  std::printf("{\"mode\": \"%s\", \"gpus\": %d, \"patients\": %d, \"slices\": %ld, \"ok\": %ld, \"seconds\": %.4f, "
              "\"slices_per_sec\": %.1f, \"startup_seconds\": %.4f, \"export\": %s}\n",
              opt.mode.c_str(), sequential ? 1 : G, P, tot.total.load(), tot.ok.load(), sec, tot.ok.load() / sec, startup,
              opt.do_export ? "true" : "false");
  if (std::getenv("NM03_DRIVER_TIMING"))
    std::fprintf(stderr, "feeder wall time per phase, summed over feeders [ms]: list %.1f, parse %.1f, read %.1f, gpu call %.1f, export submit %.1f; "
                 "export core time summed over pool threads %.1f\n",
                 tot.ns_list / 1e6, tot.ns_parse / 1e6, tot.ns_read / 1e6, tot.ns_gpu / 1e6, tot.ns_submit / 1e6, tot.ns_export / 1e6);
  nm03_shutdown(ctx);
  return 0;
}

int runTestPipeline(const Options& opt) {
  // the reference prototype through the FAST-shaped shim (test_pipeline.cpp:33-177), no window
  using namespace nm03;
  auto importer = DICOMFileImporter::create(opt.test_file);
  importer->setLoadSeries(false);
  importer->update();
  auto image = importer->getOutputData<Image>(0);
  const int width = image->getWidth(), height = image->getHeight();
  auto normalize = IntensityNormalization::create(0.5f, 2.5f, 0.0f, 10000.0f)->connect(importer);
  auto clipping = IntensityClipping::create(0.68f, 4000.0f)->connect(normalize);
  auto median = VectorMedianFilter::create(7)->connect(clipping);
  auto sharpen = ImageSharpening::create(2.0f, 0.5f, 9)->connect(median);
  std::vector<int> seeds(2 * 512);
  const int ns = nm03_make_seeds(width, height, seeds.data(), 512);
  if (ns < 0) throw Exception(nm03_last_error());
  std::vector<Vector3i> pts;
  for (int i = 0; i < ns; ++i) pts.push_back(Vector3i(seeds[2 * i], seeds[2 * i + 1], 0));
  auto rg = SeededRegionGrowing::create(0.74f, 0.91f, pts)->connect(sharpen);
  auto caster = ImageCaster::create(TYPE_UINT8)->connect(rg);
  auto erosion = Erosion::create(3)->connect(caster);
  erosion->update();
  auto dilation = Dilation::create(3)->connect(caster);
  dilation->update();
  LabelColors colors;
  colors[1] = Color::White();
  setupOutputDirectory(opt.out);
  auto rti = RenderToImage::create(Color::Black(), 512, 512);
  std::vector<std::pair<std::string, std::shared_ptr<Renderer>>> pairs = {
      {"original_image", ImageRenderer::create()->connect(importer)},
      {"preprocessed_image", ImageRenderer::create()->connect(sharpen)},
      {"segmentation", SegmentationRenderer::create(colors, 0.6f, 1.0f, 2)->connect(rg)},
      {"erosion_result", SegmentationRenderer::create(colors, 0.6f, 1.0f, 2)->connect(erosion)},
      {"final_dilated_result", SegmentationRenderer::create(colors, 0.6f, 1.0f, 2)->connect(dilation)}};
  for (auto& [name, r] : pairs) {
    rti->removeAllRenderers();
    rti->connect(r);
    rti->update();
    ImageFileExporter::create(opt.out + "/" + name + ".jpg")->connect(rti->getOutputData<Image>(0))->update();
  }
  std::cout << "test pipeline: " << width << "x" << height << ", 5 images written to " << opt.out << std::endl;
  return 0;
}

int makeDataset(const Options& opt) {
  Pool pool(std::max(0, opt.host_threads - 1));
  pool.parallel_for(opt.patients, [&](int p) {
    char name[64];
    std::snprintf(name, sizeof(name), "PGBM-%03d", p + 1);
    const fs::path dir = fs::path(opt.make_dataset) / name / "01-01-2000-synthetic" ;
    fs::create_directories(dir);
    for (int s = 0; s < opt.slices; ++s) {
      const std::vector<int16_t> px = nm03::phantom(opt.size, opt.size, p, s, opt.lowcontrast ? nm03::PHANTOM_LOWCONTRAST : nm03::PHANTOM_DEFAULT);
      char fn[64];
      std::snprintf(fn, sizeof(fn), "1-%02d.dcm", s + 1);
      nm03::dicom::save((dir / fn).string(), (const uint16_t*)px.data(), opt.size, opt.size, true, s + 1, (p & 1) == 0);
    }
  });
  std::cout << "wrote " << opt.patients << " patients x " << opt.slices << " slices " << opt.size << "x" << opt.size << " under "
            << opt.make_dataset << std::endl;
  return 0;
}

}  // namespace

int main(int argc, char** argv) {
  // Canvases, JPEG buffers and mask scratch are a few hundred KiB each and live for one slice: keep them on the heap
  // (per-thread arenas) instead of one mmap/munmap pair per allocation, which serialises the export threads in the kernel.
  mallopt(M_MMAP_THRESHOLD, 1 << 30);
  mallopt(M_TRIM_THRESHOLD, 1 << 30);
  Options opt;
  opt.data = nm03::Config::getTestDataPath() + "Brain-Tumor-Progression/T1-Post-Combined-P001-P020/";  // main_sequential.cpp:83-84
  for (int i = 1; i < argc; ++i) {
    const std::string a = argv[i];
    auto val = [&]() -> std::string { if (i + 1 >= argc) { std::cerr << "missing value for " << a << std::endl; std::exit(2); } return argv[++i]; };
    if (a == "--mode") opt.mode = val();
    else if (a == "--data") opt.data = val();
    else if (a == "--out") opt.out = val();
    else if (a == "--gpus") { const std::string v = val(); for (size_t p = 0; p < v.size();) { opt.gpus.push_back(std::atoi(v.c_str() + p)); p = v.find(',', p); if (p == std::string::npos) break; ++p; } }
    else if (a == "--batch") opt.batch = std::atoi(val().c_str());
    else if (a == "--streams") opt.streams = std::atoi(val().c_str());
    else if (a == "--host-threads") opt.host_threads = std::atoi(val().c_str());
    else if (a == "--no-export") opt.do_export = false;
    else if (a == "--dump-masks") opt.dump_masks = true;
    else if (a == "--byte-masks") opt.byte_masks = true;
    else if (a == "--cpu-dct") opt.cpu_dct = true;
    else if (a == "--gpu-seg") opt.gpu_seg = true;
    else if (a == "--test-pipeline") opt.test_file = val();
    else if (a == "--make-dataset") opt.make_dataset = val();
    else if (a == "--patients") opt.patients = std::atoi(val().c_str());
    else if (a == "--slices") opt.slices = std::atoi(val().c_str());
    else if (a == "--size") opt.size = std::atoi(val().c_str());
    else if (a == "--lowcontrast") opt.lowcontrast = true;
    else { std::cerr << "unknown option " << a << std::endl; return 2; }
  }
  if (opt.host_threads <= 0) opt.host_threads = std::max(1u, std::thread::hardware_concurrency());

  if (opt.mode == "sequential" && opt.out == "../out-parallel") opt.out = "../out-sequential";
  try {
    if (!opt.make_dataset.empty()) return makeDataset(opt);
    if (!opt.test_file.empty()) { if (opt.out == "../out-parallel") opt.out = "../out-test"; return runTestPipeline(opt); }
    return runBatch(opt);
  } catch (const std::exception& e) {
    std::cerr << "Fatal error: " << e.what() << std::endl;
    return 1;
  }
}
