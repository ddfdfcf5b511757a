// slc_hostxfer.cu -- moving the lent host arrays of upload / download / compress at link speed.
//
// The reference's callers own plain Storable vectors (Data/Matrix/Sparse/Foreign.hs:33-55): pageable
// memory, often fresh from the allocator.  Measured on the B200 box (scripts/r3/probe_host_link.cu,
// profiles/r3_probe_host_link.txt, 512 MiB):
//     pinned H2D / D2H                          55.6 / 54.4 GB/s
//     pageable H2D (the driver's own staging)   11.3 GB/s
//     pageable D2H, pages touched before        19.4 GB/s
//     pageable D2H, pages never touched          2.1 GB/s      <- what a fresh result buffer gets
//     memcpy pageable -> pinned, 1 / 4 threads  14.9 / 50.0 GB/s
// so a cudaMemcpyAsync on the caller's pointer leaves 5x (in) to 25x (out) on the table, and round 2's
// mm / transpose / compress spent 58-346 ms around 0.5-1.4 ms of kernels.  Here a large transfer whose host
// side is not pinned goes through a ring of pinned buffers: a few host threads copy a chunk between the
// caller's memory and a ring slot while the DMA engine moves the previous chunks.  The threads that copy
// OUT also take the page faults of an untouched destination in parallel (the kernel zeroes fresh pages at
// about 2.3 GB/s per thread), after the range has been offered to the kernel as huge pages.
//
// One ring per process (4 x 16 MiB, allocated at the first large pageable transfer), one transfer at a
// time (a mutex: transfers are bandwidth-bound, two at once would share the same link).  When a transfer
// returns the ring is idle again, so a slot never carries an event of another device into the next call.
#include "slc_common.cuh"

#include <sys/mman.h>

#include <algorithm>
#include <thread>

namespace slc {

namespace {

constexpr int XFER_SLOTS = 4;
constexpr size_t XFER_CHUNK = (size_t)16 << 20;
constexpr size_t XFER_MIN_BYTES = (size_t)4 << 20;  // below this the driver's path costs less than the set-up

struct Ring {
  std::mutex mu;
  bool tried = false;
  char *slot[XFER_SLOTS] = {};
};
Ring g_ring;

// per-device events of the slots (an event can only be recorded on a stream of its own device)
struct RingEvents {
  bool made = false;
  cudaEvent_t ev[XFER_SLOTS];
};
RingEvents g_ring_events[MAX_DEVICES];

int copy_threads() {
  static const int n = [] {
    const int env = env_int("SLC_COPY_THREADS", 0);
    if (env > 0) return std::min(env, 32);
    const unsigned hw = std::thread::hardware_concurrency();
    return (int)std::max(1u, std::min(8u, hw / 2));
  }();
  return n;
}

void parallel_memcpy(char *dst, const char *src, size_t bytes) {
  const int nt = copy_threads();
  if (nt <= 1 || bytes < ((size_t)1 << 20)) {
    memcpy(dst, src, bytes);
    return;
  }
  const size_t per = ((bytes + nt - 1) / nt + 4095) & ~(size_t)4095;
  std::thread th[32];
  int started = 0;
  for (int t = 1; t < nt; ++t) {
    const size_t off = (size_t)t * per;
    if (off >= bytes) break;
    const size_t len = std::min(per, bytes - off);
    th[started++] = std::thread([=] { memcpy(dst + off, src + off, len); });
  }
  memcpy(dst, src, std::min(per, bytes));
  for (int t = 0; t < started; ++t) th[t].join();
}

bool is_pinned(const void *p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    (void)cudaGetLastError();
    return false;
  }
  return a.type == cudaMemoryTypeHost || a.type == cudaMemoryTypeManaged;
}

// the ring and this device's events, or false when they cannot be had (the caller falls back to the plain copy)
bool ring_ready(cudaEvent_t **ev) {
  if (!g_ring.tried) {
    g_ring.tried = true;
    for (int i = 0; i < XFER_SLOTS; ++i)
      if (cudaHostAlloc((void **)&g_ring.slot[i], XFER_CHUNK, cudaHostAllocPortable) != cudaSuccess) {
        (void)cudaGetLastError();
        for (int j = 0; j < i; ++j) cudaFreeHost(g_ring.slot[j]);
        for (auto &s : g_ring.slot) s = nullptr;
        break;
      }
  }
  if (!g_ring.slot[0]) return false;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAX_DEVICES) return false;
  RingEvents &re = g_ring_events[dev];
  if (!re.made) {
    for (int i = 0; i < XFER_SLOTS; ++i)
      if (cudaEventCreateWithFlags(&re.ev[i], cudaEventDisableTiming) != cudaSuccess) {
        (void)cudaGetLastError();
        for (int j = 0; j < i; ++j) cudaEventDestroy(re.ev[j]);
        return false;
      }
    re.made = true;
  }
  *ev = re.ev;
  return true;
}

bool staging_wanted(const void *host, size_t bytes) {
  static const int on = env_int("SLC_HOST_STAGING", 1);  // 0: always hand the caller's pointer to cudaMemcpyAsync
  return on != 0 && bytes >= XFER_MIN_BYTES && !is_pinned(host);
}

}  // namespace

bool host_needs_staging(const void *host, size_t bytes) { return staging_wanted(host, bytes); }

// A block of `height` columns of `width` bytes each (pitches in bytes), cut into pieces of at most one ring
// slot: piece q is bytes [off, off + len) of column col.
struct Pieces {
  size_t width, height, per_col, total;
  Pieces(size_t w, size_t h) : width(w), height(h), per_col((w + XFER_CHUNK - 1) / XFER_CHUNK), total(per_col * h) {}
  void at(size_t q, size_t *col, size_t *off, size_t *len) const {
    *col = q / per_col;
    *off = (q % per_col) * XFER_CHUNK;
    *len = std::min(XFER_CHUNK, width - *off);
  }
};

// dst_dev <- src_host, `height` columns of `width` bytes.  On return the source has been read (the caller
// may reuse it); the device side completes in stream order.  The pieces of all columns form one pipeline:
// the host threads fill slot q+1 while the DMA engine drains slot q.
int host_to_dev_2d(void *dst_dev, size_t dpitch, const void *src_host, size_t spitch, size_t width, size_t height,
                   cudaStream_t s) {
  if (width == 0 || height == 0) return SLC_OK;
  if (staging_wanted(src_host, width)) {
    std::lock_guard<std::mutex> lock(g_ring.mu);
    cudaEvent_t *ev = nullptr;
    if (ring_ready(&ev)) {
      const char *src = static_cast<const char *>(src_host);
      char *dst = static_cast<char *>(dst_dev);
      const Pieces pc(width, height);
      for (size_t q = 0; q < pc.total; ++q) {
        size_t col, off, len;
        pc.at(q, &col, &off, &len);
        const int k = (int)(q % XFER_SLOTS);
        if (q >= (size_t)XFER_SLOTS) SLC_CUDA(cudaEventSynchronize(ev[k]));  // the DMA that last read this slot
        parallel_memcpy(g_ring.slot[k], src + col * spitch + off, len);
        SLC_CUDA(cudaMemcpyAsync(dst + col * dpitch + off, g_ring.slot[k], len, cudaMemcpyHostToDevice, s));
        SLC_CUDA(cudaEventRecord(ev[k], s));
      }
      for (size_t k = 0; k < std::min<size_t>(pc.total, XFER_SLOTS); ++k) SLC_CUDA(cudaEventSynchronize(ev[k]));  // ring idle again
      return SLC_OK;
    }
  }
  if (height == 1) SLC_CUDA(cudaMemcpyAsync(dst_dev, src_host, width, cudaMemcpyHostToDevice, s));
  else SLC_CUDA(cudaMemcpy2DAsync(dst_dev, dpitch, src_host, spitch, width, height, cudaMemcpyHostToDevice, s));
  return SLC_OK;
}

// dst_host <- src_dev, in stream order behind whatever produces src_dev.  The destination is complete after
// the caller's next synchronisation of `s` (the staged path has already waited for it).
int dev_to_host_2d(void *dst_host, size_t dpitch, const void *src_dev, size_t spitch, size_t width, size_t height,
                   cudaStream_t s) {
  if (width == 0 || height == 0) return SLC_OK;
  if (staging_wanted(dst_host, width)) {
    std::lock_guard<std::mutex> lock(g_ring.mu);
    cudaEvent_t *ev = nullptr;
    if (ring_ready(&ev)) {
      char *dst = static_cast<char *>(dst_host);
      const char *src = static_cast<const char *>(src_dev);
      for (size_t c = 0; c < height; ++c) {  // a hint only: fresh anonymous memory is then faulted in 2 MiB at a time where the system allows it
        const uintptr_t b = (uintptr_t)(dst + c * dpitch);
        const uintptr_t lo = (b + 4095) & ~(uintptr_t)4095, hi = (b + width) & ~(uintptr_t)4095;
        if (hi > lo) (void)madvise((void *)lo, hi - lo, MADV_HUGEPAGE);
      }
      const Pieces pc(width, height);
      for (size_t q = 0; q < pc.total + XFER_SLOTS; ++q) {
        if (q >= (size_t)XFER_SLOTS) {  // piece j has landed in its slot: copy it out, which frees the slot for piece q
          const size_t j = q - XFER_SLOTS;
          size_t col, off, len;
          pc.at(j, &col, &off, &len);
          SLC_CUDA(cudaEventSynchronize(ev[j % XFER_SLOTS]));
          parallel_memcpy(dst + col * dpitch + off, g_ring.slot[j % XFER_SLOTS], len);
        }
        if (q < pc.total) {
          size_t col, off, len;
          pc.at(q, &col, &off, &len);
          SLC_CUDA(cudaMemcpyAsync(g_ring.slot[q % XFER_SLOTS], src + col * spitch + off, len, cudaMemcpyDeviceToHost, s));
          SLC_CUDA(cudaEventRecord(ev[q % XFER_SLOTS], s));
        }
      }
      return SLC_OK;
    }
  }
  if (height == 1) SLC_CUDA(cudaMemcpyAsync(dst_host, src_dev, width, cudaMemcpyDeviceToHost, s));
  else SLC_CUDA(cudaMemcpy2DAsync(dst_host, dpitch, src_dev, spitch, width, height, cudaMemcpyDeviceToHost, s));
  return SLC_OK;
}

int host_to_dev(void *dst_dev, const void *src_host, size_t bytes, cudaStream_t s) {
  return host_to_dev_2d(dst_dev, bytes, src_host, bytes, bytes, 1, s);
}
int dev_to_host(void *dst_host, const void *src_dev, size_t bytes, cudaStream_t s) {
  return dev_to_host_2d(dst_host, bytes, src_dev, bytes, bytes, 1, s);
}

}  // namespace slc
