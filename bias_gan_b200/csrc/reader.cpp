// .npy sample reader behind the reference's numpy_reader interface
// (src/numpy_reader/cpp/numpy_reader.cpp:58-185 parse, :342-380 init/finalize, :403-497 reads,
//  src/numpy_reader/cpp/thread_pool.cpp:24-155 worker pool).
//
// Design (not a port): a small persistent worker pool pulls chunk jobs; each worker preads its
// chunk either straight into the caller's host buffer or into a pinned staging ring and then
// issues its own cudaMemcpyAsync, so disk reads of later chunks overlap the H2D DMA of earlier
// ones ("GDS-style" chunked direct placement without needing nvidia-fs on the box).  The caller
// owns the destination buffer: no IO-buffer clone as in numpy_reader.cpp:474,496.
// Short files raise "file corruption" instead of spinning (reference bug, numpy_reader.cpp:403-431).
#include <errno.h>
#include <fcntl.h>
#include <string.h>
#include <sys/stat.h>
#include <unistd.h>

#include <atomic>
#include <condition_variable>
#include <deque>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

#include "bgr_internal.h"

namespace bgr {

// ------------------------------------------------------------------------------ worker pool
class Pool {
 public:
  Pool(int nthreads, int device) : device_(device) {
    if (nthreads < 1) nthreads = 1;
    for (int i = 0; i < nthreads; ++i) threads_.emplace_back([this] { loop(); });
  }
  ~Pool() {
    {
      std::lock_guard<std::mutex> g(m_);
      stop_ = true;
    }
    cv_.notify_all();
    for (auto& t : threads_) t.join();
  }
  int size() const { return int(threads_.size()); }
  // Run all jobs; a job returns 0 or an error code and fills msg.  Returns the first failure.
  int run(std::vector<std::function<int(std::string&)>>& jobs, std::string& msg) {
    if (jobs.empty()) return BGR_OK;
    {
      std::lock_guard<std::mutex> g(m_);
      err_code_ = BGR_OK;
      err_msg_.clear();
      pending_ = int(jobs.size());
      for (auto& j : jobs) q_.push_back(&j);
    }
    cv_.notify_all();
    std::unique_lock<std::mutex> lk(m_);
    done_cv_.wait(lk, [this] { return pending_ == 0; });
    msg = err_msg_;
    return err_code_;
  }

 private:
  void loop() {
    if (device_ >= 0) cudaSetDevice(device_);
    for (;;) {
      std::function<int(std::string&)>* job = nullptr;
      {
        std::unique_lock<std::mutex> lk(m_);
        cv_.wait(lk, [this] { return stop_ || !q_.empty(); });
        if (stop_ && q_.empty()) return;
        job = q_.front();
        q_.pop_front();
      }
      std::string msg;
      int rc = BGR_ERR_RUNTIME;
      try {
        rc = (*job)(msg);
      } catch (const std::exception& e) {
        msg = e.what();
      }
      {
        std::lock_guard<std::mutex> g(m_);
        if (rc != BGR_OK && err_code_ == BGR_OK) {
          err_code_ = rc;
          err_msg_ = msg;
        }
        if (--pending_ == 0) done_cv_.notify_all();
      }
    }
  }
  int device_;
  std::vector<std::thread> threads_;
  std::deque<std::function<int(std::string&)>*> q_;
  std::mutex m_;
  std::condition_variable cv_, done_cv_;
  bool stop_ = false;
  int pending_ = 0;
  int err_code_ = BGR_OK;
  std::string err_msg_;
};

// ------------------------------------------------------------------------------ reader state
struct Reader {
  bool split_axis = false;
  int device = -1;
  int inter = 1, intra = 1;
  Pool* pool = nullptr;
  bool parsed = false, fortran = false;
  int dtype = -1, typesize = 0;
  std::vector<int64_t> shape, stride;
  int64_t numsample = 1, numelem = 0, offset = 0, batchsize = 1;
  int fd = -1;
  // pinned staging ring for the device path: two slots alternate between calls
  unsigned char* pinned[2] = {nullptr, nullptr};
  size_t pinned_bytes[2] = {0, 0};
  cudaEvent_t slot_free[2] = {nullptr, nullptr};
  int next_slot = 0;
  int64_t bytes_read = 0;  // payload bytes delivered since creation (for reader GB/s)

  ~Reader() {
    if (fd >= 0) close(fd);
    delete pool;
    for (int i = 0; i < 2; ++i) {
      if (slot_free[i]) {
        cudaEventSynchronize(slot_free[i]);
        cudaEventDestroy(slot_free[i]);
      }
      if (pinned[i]) cudaFreeHost(pinned[i]);
    }
  }
  void rebuild_pool() {
    delete pool;
    pool = new Pool(inter * intra, device);
  }
};

static bool find_value(const std::string& h, const char* key, size_t* pos) {
  size_t k = h.find(key);
  if (k == std::string::npos) return false;
  k = h.find(':', k + strlen(key));
  if (k == std::string::npos) return false;
  ++k;
  while (k < h.size() && h[k] == ' ') ++k;
  *pos = k;
  return k < h.size();
}

static int parse_file(Reader* r, const char* filename) {
  int fd = open(filename, O_RDONLY);
  if (fd < 0)
    return fail(BGR_ERR_RUNTIME,
                std::string("NumpyReader: failed to open file ") + filename + ": " + strerror(errno));
  unsigned char magic[12];
  ssize_t n = pread(fd, magic, 12, 0);
  if (n < 10) {
    close(fd);
    if (n >= 6 && memcmp(magic, "\x93NUMPY", 6) == 0)
      return fail(BGR_ERR_RUNTIME, "NumpyReader: could not read header, ill formatted or corrupt.");
    return fail(BGR_ERR_RUNTIME, "NumpyReader: not a numpy file.");
  }
  if (memcmp(magic, "\x93NUMPY", 6) != 0) {
    close(fd);
    return fail(BGR_ERR_RUNTIME, "NumpyReader: not a numpy file.");
  }
  const int major = magic[6];
  size_t header_len, header_off;
  if (major == 1) {
    header_len = size_t(magic[8]) | (size_t(magic[9]) << 8);
    header_off = 10;
  } else if ((major == 2 || major == 3) && n >= 12) {
    header_len = size_t(magic[8]) | (size_t(magic[9]) << 8) | (size_t(magic[10]) << 16) |
                 (size_t(magic[11]) << 24);
    header_off = 12;
  } else {
    close(fd);
    return fail(BGR_ERR_RUNTIME,
                "NumpyReader: cannot parse header, ill formatted or corrupt (bad version bytes).");
  }
  struct stat stt;
  fstat(fd, &stt);
  if (header_len == 0 || off_t(header_off + header_len) > stt.st_size) {
    close(fd);
    return fail(BGR_ERR_RUNTIME, "NumpyReader: cannot parse header, ill formatted or corrupt.");
  }
  std::string h(header_len, '\0');
  n = pread(fd, &h[0], header_len, header_off);
  close(fd);
  if (n != ssize_t(header_len))
    return fail(BGR_ERR_RUNTIME, "NumpyReader: cannot parse header, ill formatted or corrupt.");

  size_t p_descr, p_order, p_shape;
  if (h.empty() || h[0] != '{' || !find_value(h, "'descr'", &p_descr) ||
      !find_value(h, "'fortran_order'", &p_order) || !find_value(h, "'shape'", &p_shape) ||
      h[p_descr] != '\'' || h[p_shape] != '(')
    return fail(BGR_ERR_RUNTIME, "NumpyReader: cannot parse header, ill formatted or corrupt.");
  size_t q = h.find('\'', p_descr + 1);
  size_t s_end = h.find(')', p_shape);
  if (q == std::string::npos || s_end == std::string::npos || q - p_descr < 3)
    return fail(BGR_ERR_RUNTIME, "NumpyReader: cannot parse header, ill formatted or corrupt.");
  const std::string descr = h.substr(p_descr + 1, q - p_descr - 1);
  const char endian = descr[0];
  if (!(endian == '<' || endian == '|' || endian == '='))
    return fail(BGR_ERR_RUNTIME,
                "NumpyReader: the specified file is in big endian. This is currently not supported.");
  const std::string tid = descr.substr(1);
  int dtype, tsize;
  if (tid == "i8") { dtype = BGR_I64; tsize = 8; }
  else if (tid == "i4") { dtype = BGR_I32; tsize = 4; }
  else if (tid == "f8") { dtype = BGR_F64; tsize = 8; }
  else if (tid == "f4") { dtype = BGR_F32; tsize = 4; }
  else return fail(BGR_ERR_RUNTIME, "NumpyReader: unsupported datatype " + descr + ".");
  bool fortran;
  if (h.compare(p_order, 4, "True") == 0) fortran = true;
  else if (h.compare(p_order, 5, "False") == 0) fortran = false;
  else return fail(BGR_ERR_RUNTIME, "NumpyReader: cannot parse header, ill formatted or corrupt.");

  std::vector<int64_t> shape;
  {
    size_t i = p_shape + 1;
    while (i < s_end) {
      while (i < s_end && (h[i] == ' ' || h[i] == ',')) ++i;
      if (i >= s_end) break;
      if (h[i] < '0' || h[i] > '9')
        return fail(BGR_ERR_RUNTIME, "NumpyReader: cannot parse header, ill formatted or corrupt.");
      int64_t v = 0;
      while (i < s_end && h[i] >= '0' && h[i] <= '9') v = v * 10 + (h[i++] - '0');
      shape.push_back(v);
    }
  }
  if (shape.empty()) shape.push_back(1);  // scalar "()" -> one element (numpy_reader.cpp:145-146)

  int64_t numsample = 1;
  if (r->split_axis) {
    if (fortran)
      return fail(BGR_ERR_RUNTIME,
                  "NumpyReader: reading column-major arrays (Fortran order) is currently only "
                  "supported if the split_axis option is false.");
    numsample = shape[0];
    shape.erase(shape.begin());
    if (shape.empty()) shape.push_back(1);
  }
  r->dtype = dtype;
  r->typesize = tsize;
  r->fortran = fortran;
  r->shape = shape;
  r->numsample = numsample;
  r->offset = int64_t(header_off + header_len);
  r->numelem = 1;
  for (int64_t d : shape) r->numelem *= d;
  r->stride.assign(shape.size(), 1);
  if (!fortran) {
    for (int i = int(shape.size()) - 2; i >= 0; --i) r->stride[i] = r->stride[i + 1] * shape[i + 1];
  } else {
    for (size_t i = 1; i < shape.size(); ++i) r->stride[i] = r->stride[i - 1] * shape[i - 1];
  }
  r->parsed = true;
  return BGR_OK;
}

// pread exactly `size` bytes at `off`; a short file is corruption, never a spin.
static int read_fully(int fd, unsigned char* dst, int64_t size, int64_t off, std::string& msg) {
  while (size > 0) {
    ssize_t n = pread(fd, dst, size_t(size), off_t(off));
    if (n < 0) {
      if (errno == EINTR) continue;
      msg = std::string("NumpyReader: read failed: ") + strerror(errno);
      return BGR_ERR_RUNTIME;
    }
    if (n == 0) {
      msg = "NumpyReader: file corruption, payload is shorter than the header promises.";
      return BGR_ERR_INDEX;
    }
    dst += n;
    off += n;
    size -= n;
  }
  return BGR_OK;
}

}  // namespace bgr

using namespace bgr;

extern "C" int bgr_reader_create(int split_axis, int device, void** handle) {
  if (!handle) return fail(BGR_ERR_ARG, "bgr_reader_create: null handle");
  if (device >= 0) {
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || device >= count)
      return fail(BGR_ERR_CUDA, "NumpyReader: CUDA device " + std::to_string(device) +
                                    " is not available (no CPU fallback for the device path).");
  }
  Reader* r = new Reader();
  r->split_axis = split_axis != 0;
  r->device = device;
  r->rebuild_pool();
  *handle = r;
  return BGR_OK;
}

extern "C" int bgr_reader_destroy(void* handle) {
  delete static_cast<Reader*>(handle);
  return BGR_OK;
}

extern "C" int bgr_reader_parse(void* handle, const char* filename) {
  return parse_file(static_cast<Reader*>(handle), filename);
}

extern "C" int bgr_reader_init_file(void* handle, const char* filename) {
  Reader* r = static_cast<Reader*>(handle);
  if (r->fd >= 0) close(r->fd);
  r->fd = open(filename, O_RDONLY);
  if (r->fd < 0)
    return fail(BGR_ERR_RUNTIME,
                std::string("NumpyReader: failed to open file ") + filename + ": " + strerror(errno));
  posix_fadvise(r->fd, 0, 0, POSIX_FADV_SEQUENTIAL);
  // The reference re-uses the geometry of the file that was parsed for every file it opens afterwards
  // (gpsro_dataset.py:133-148).  A LARGER file than the parsed header describes is a file of another shape: its
  // bytes would be delivered silently mis-shaped, so refuse it.  (A shorter one keeps the reference's
  // behaviour: IndexError("... file corruption ...") from the read.)
  if (r->parsed) {
    struct stat sb;
    const int64_t expect = r->offset + r->numsample * r->numelem * int64_t(r->typesize);
    if (fstat(r->fd, &sb) == 0 && int64_t(sb.st_size) > expect + 4096) {
      close(r->fd);
      r->fd = -1;
      return fail(BGR_ERR_RUNTIME, std::string("NumpyReader: ") + filename + " holds " +
                                       std::to_string(int64_t(sb.st_size)) + " bytes, more than the parsed header describes (" +
                                       std::to_string(expect) + "): call parse() for files of another shape");
    }
  }
  return BGR_OK;
}

extern "C" int bgr_reader_finalize_file(void* handle) {
  Reader* r = static_cast<Reader*>(handle);
  if (r->fd >= 0) close(r->fd);
  r->fd = -1;
  return BGR_OK;
}

extern "C" int bgr_reader_set_batchsize(void* handle, int64_t batch_size) {
  Reader* r = static_cast<Reader*>(handle);
  if (batch_size <= 0 || batch_size > r->numsample)
    return fail(BGR_ERR_INDEX,
                "NumpyReader: the batch size has to be a positive number and must not be bigger "
                "than the total number of samples.");
  if (!r->split_axis && batch_size > 1)
    return fail(BGR_ERR_INDEX,
                "NumpyReader: in order to use batching, you must have more than one sample per "
                "file. Otherwise batch externally.");
  r->batchsize = batch_size;
  return BGR_OK;
}

extern "C" int bgr_reader_set_threads(void* handle, int inter, int intra) {
  Reader* r = static_cast<Reader*>(handle);
  if (inter < 1 || intra < 1) return fail(BGR_ERR_ARG, "NumpyReader: thread counts must be >= 1");
  if (inter != r->inter || intra != r->intra) {
    r->inter = inter;
    r->intra = intra;
    r->rebuild_pool();
  }
  return BGR_OK;
}

extern "C" int bgr_reader_info(void* handle, int* ndim, int64_t* shape, int64_t* stride, int* dtype,
                               int64_t* num_samples, int* fortran_order, int64_t* payload_offset,
                               int64_t* batchsize, int* inter, int* intra) {
  Reader* r = static_cast<Reader*>(handle);
  if (ndim) *ndim = int(r->shape.size());
  for (size_t i = 0; i < r->shape.size() && i < 8; ++i) {
    if (shape) shape[i] = r->shape[i];
    if (stride) stride[i] = r->stride[i];
  }
  if (dtype) *dtype = r->dtype;
  if (num_samples) *num_samples = r->numsample;
  if (fortran_order) *fortran_order = r->fortran ? 1 : 0;
  if (payload_offset) *payload_offset = r->offset;
  if (batchsize) *batchsize = r->batchsize;
  if (inter) *inter = r->inter;
  if (intra) *intra = r->intra;
  return BGR_OK;
}

// Read n samples (ids wrap like numpy_reader.cpp:438) into dst, sample-major.  dst_device < 0:
// dst is host memory.  Otherwise dst is device memory on cuda:dst_device; chunks are staged
// through pinned memory and copied on `stream`; the call returns when the copies have been
// enqueued and the pinned slot is protected by an event (stream-ordered result).
extern "C" int bgr_reader_read(void* handle, const int64_t* ids, int n, void* dst, int dst_device,
                               void* stream) {
  Reader* r = static_cast<Reader*>(handle);
  if (!r->parsed) return fail(BGR_ERR_RUNTIME, "NumpyReader: parse() has to be called before reading.");
  if (r->fd < 0) return fail(BGR_ERR_RUNTIME, "NumpyReader: no file open, call init_file() first.");
  if (n != r->batchsize)
    return fail(BGR_ERR_INDEX, "NumpyReader: number of requested elements does not match the batch size.");
  const int64_t bsize = r->numelem * r->typesize;
  const int64_t total = bsize * n;
  unsigned char* stage = nullptr;
  int slot = 0;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (dst_device >= 0) {
    BGR_CUDA_OK(cudaSetDevice(dst_device));
    slot = r->next_slot;
    r->next_slot ^= 1;
    if (!r->slot_free[slot]) BGR_CUDA_OK(cudaEventCreateWithFlags(&r->slot_free[slot], cudaEventDisableTiming));
    else BGR_CUDA_OK(cudaEventSynchronize(r->slot_free[slot]));
    if (r->pinned_bytes[slot] < size_t(total)) {
      if (r->pinned[slot]) cudaFreeHost(r->pinned[slot]);
      r->pinned[slot] = nullptr;
      BGR_CUDA_OK(cudaHostAlloc(reinterpret_cast<void**>(&r->pinned[slot]), size_t(total), cudaHostAllocDefault));
      r->pinned_bytes[slot] = size_t(total);
    }
    stage = r->pinned[slot];
  }
  // chunking: up to `intra` chunks per sample, 4 KiB aligned when the sample is large
  int64_t nth = (bsize / r->intra > 1) ? r->intra : 1;
  int64_t chunk = (bsize + nth - 1) / nth;
  if (chunk > 65536) chunk = (chunk + 4095) & ~int64_t(4095);
  std::vector<std::function<int(std::string&)>> jobs;
  const int fd = r->fd;
  for (int b = 0; b < n; ++b) {
    const int64_t id = ((ids[b] % r->numsample) + r->numsample) % r->numsample;
    const int64_t src0 = r->offset + id * bsize;
    const int64_t dst0 = int64_t(b) * bsize;
    for (int64_t off = 0; off < bsize; off += chunk) {
      const int64_t sz = std::min(chunk, bsize - off);
      unsigned char* hdst = (stage ? stage : static_cast<unsigned char*>(dst)) + dst0 + off;
      unsigned char* ddst = static_cast<unsigned char*>(dst) + dst0 + off;
      const bool to_dev = stage != nullptr;
      jobs.emplace_back([=](std::string& msg) -> int {
        int rc = read_fully(fd, hdst, sz, src0 + off, msg);
        if (rc != BGR_OK) return rc;
        if (to_dev) {
          cudaError_t e = cudaMemcpyAsync(ddst, hdst, size_t(sz), cudaMemcpyHostToDevice, st);
          if (e != cudaSuccess) {
            msg = std::string("NumpyReader: cudaMemcpyAsync failed: ") + cudaGetErrorString(e);
            return BGR_ERR_CUDA;
          }
        }
        return BGR_OK;
      });
    }
  }
  std::string msg;
  int rc = r->pool->run(jobs, msg);
  if (dst_device >= 0) cudaEventRecord(r->slot_free[slot], st);
  if (rc != BGR_OK) return fail(rc, msg);
  r->bytes_read += total;
  return BGR_OK;
}

extern "C" int64_t bgr_reader_bytes_read(void* handle) { return static_cast<Reader*>(handle)->bytes_read; }

// All-pairs peer access (numpy_reader.cpp:17-56).  Already-enabled pairs are not an error.
extern "C" int bgr_reader_enable_p2p(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) return fail(BGR_ERR_CUDA, "NumpyReader: no CUDA devices");
  int cur = 0;
  cudaGetDevice(&cur);
  for (int i = 0; i < n; ++i) {
    cudaSetDevice(i);
    for (int j = 0; j < n; ++j) {
      if (i == j) continue;
      int can = 0;
      cudaDeviceCanAccessPeer(&can, i, j);
      if (!can) continue;
      cudaError_t e = cudaDeviceEnablePeerAccess(j, 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled)
        return fail(BGR_ERR_CUDA, std::string("NumpyReader: enable_p2p: ") + cudaGetErrorString(e));
      cudaGetLastError();
    }
  }
  cudaSetDevice(cur);
  return BGR_OK;
}

extern "C" int bgr_reader_disable_p2p(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) return fail(BGR_ERR_CUDA, "NumpyReader: no CUDA devices");
  int cur = 0;
  cudaGetDevice(&cur);
  for (int i = 0; i < n; ++i) {
    cudaSetDevice(i);
    for (int j = 0; j < n; ++j) {
      if (i == j) continue;
      cudaError_t e = cudaDeviceDisablePeerAccess(j);
      (void)e;
      cudaGetLastError();
    }
  }
  cudaSetDevice(cur);
  return BGR_OK;
}
