// hostio.cu -- host <-> device transfers of the C ABI (upload/download of particles and
// fields).  The caller's buffers are ordinary pageable memory (Fortran arrays, numpy);
// a cudaMemcpy from pageable memory runs at ~10 GB/s because the driver stages it through
// one small pinned buffer on one thread.  Here the staging is explicit: a ring of pinned
// buffers, filled (or drained) by several host threads while the DMA of the previous chunk
// is in flight, so the transfer runs at the PCIe rate.  Everything is issued on the
// context's stream; the host buffer is fully consumed (or filled) when the call returns.
#include "ctx.cuh"
#include <algorithm>
#include <cstring>
#include <functional>
#include <thread>
#include <vector>

namespace eb {

static void parallel_memcpy(void* dst, const void* src, size_t bytes, int nthreads) {
  if (bytes < (size_t)(8 << 20) || nthreads <= 1) { memcpy(dst, src, bytes); return; }
  std::vector<std::thread> th;
  const size_t per = ((bytes / nthreads) + 4095) & ~(size_t)4095;
  for (int t = 1; t < nthreads; ++t) {
    const size_t o = (size_t)t * per;
    if (o >= bytes) break;
    const size_t m = std::min(per, bytes - o);
    th.emplace_back([=] { memcpy((char*)dst + o, (const char*)src + o, m); });
  }
  memcpy(dst, src, std::min(per, bytes));
  for (std::thread& t : th) t.join();
}

int32_t ensure_pinned(eb200_ctx* c) {
  if (c->pin[0]) return 0;
  if (c->host_threads <= 0) {
    const unsigned hc = std::thread::hardware_concurrency();
    c->host_threads = (int)std::max(1u, std::min(12u, hc > 4 ? hc - 2 : 1u));   // 12 saturate PCIe 5 x16 (measured)
  }
  for (int b = 0; b < eb200_ctx::NPIN; ++b) {
    EB_CUDA(c, cudaHostAlloc(&c->pin[b], c->pin_bytes, cudaHostAllocDefault));
    EB_CUDA(c, cudaEventCreateWithFlags(&c->pin_ev[b], cudaEventDisableTiming));
    c->pin_busy[b] = false;
  }
  return 0;
}

void free_pinned(eb200_ctx* c) {
  for (int b = 0; b < eb200_ctx::NPIN; ++b) {
    if (c->pin[b]) { cudaFreeHost(c->pin[b]); cudaEventDestroy(c->pin_ev[b]); }
    c->pin[b] = nullptr;
  }
}

// acquire the next pinned buffer of the ring (waits for the DMA that last used it)
static int32_t pin_acquire(eb200_ctx* c, int& b) {
  b = c->pin_next;
  c->pin_next = (c->pin_next + 1) % eb200_ctx::NPIN;
  if (c->pin_busy[b]) { EB_CUDA(c, cudaEventSynchronize(c->pin_ev[b])); c->pin_busy[b] = false; }
  return 0;
}

// page-locked (eb200_host_alloc / cudaHostAlloc / cudaHostRegister) buffers need no staging
static bool is_pinned(const void* host) {
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, host) != cudaSuccess) { cudaGetLastError(); return false; }
  return at.type == cudaMemoryTypeHost;
}

// host -> device.  The transfer is cut into chunks of whole granules (<= one pinned buffer);
// dst(o, m) names the device address chunk [o, o + m) is copied to and after(o, m) may enqueue
// a kernel that consumes it (both stream-ordered, so one device staging area can be reused).
int32_t staged_h2d(eb200_ctx* c, const void* host, size_t bytes, size_t granule,
                   const std::function<void*(size_t, size_t)>& dst, const std::function<int32_t(size_t, size_t)>& after) {
  if (bytes == 0) return 0;
  EB_TRY(ensure_pinned(c));
  const size_t chunk = std::max<size_t>(1, c->pin_bytes / granule) * granule;
  if (chunk > c->pin_bytes) return fail(c, EB200_E_ARG, "transfer granule larger than the pinned staging buffer");
  if (is_pinned(host)) {          // the copy engine reads the caller's buffer directly
    for (size_t done = 0; done < bytes; done += chunk) {
      const size_t m = std::min(chunk, bytes - done);
      EB_CUDA(c, cudaMemcpyAsync(dst(done, m), (const char*)host + done, m, cudaMemcpyHostToDevice, c->stream));
      if (after) EB_TRY(after(done, m));
    }
    return 0;
  }
  for (size_t done = 0; done < bytes; done += chunk) {
    const size_t m = std::min(chunk, bytes - done);
    int b;
    EB_TRY(pin_acquire(c, b));
    parallel_memcpy(c->pin[b], (const char*)host + done, m, c->host_threads);
    EB_CUDA(c, cudaMemcpyAsync(dst(done, m), c->pin[b], m, cudaMemcpyHostToDevice, c->stream));
    EB_CUDA(c, cudaEventRecord(c->pin_ev[b], c->stream));
    c->pin_busy[b] = true;
    if (after) EB_TRY(after(done, m));
  }
  return 0;
}

// device -> host: src(o, m) enqueues whatever produces chunk [o, o + m) and returns its device
// address; the copies are issued ahead (as many as the ring holds) and drained in order
int32_t staged_d2h(eb200_ctx* c, void* host, size_t bytes, size_t granule,
                   const std::function<const void*(size_t, size_t)>& src) {
  if (bytes == 0) return 0;
  EB_TRY(ensure_pinned(c));
  const size_t chunk = std::max<size_t>(1, c->pin_bytes / granule) * granule;
  if (chunk > c->pin_bytes) return fail(c, EB200_E_ARG, "transfer granule larger than the pinned staging buffer");
  const size_t nchunk = (bytes + chunk - 1) / chunk;
  if (is_pinned(host)) {          // ... and writes it directly
    for (size_t i = 0; i < nchunk; ++i) {
      const size_t o = i * chunk, m = std::min(chunk, bytes - o);
      const void* d = src(o, m);
      if (!d) return EB200_E_CUDA;
      EB_CUDA(c, cudaMemcpyAsync((char*)host + o, d, m, cudaMemcpyDeviceToHost, c->stream));
    }
    return 0;
  }
  constexpr int NP = eb200_ctx::NPIN;
  int slot[NP];
  size_t issued = 0;
  for (size_t i = 0; i < nchunk; ++i) {
    while (issued < nchunk && issued < i + NP) {
      const size_t o = issued * chunk, m = std::min(chunk, bytes - o);
      int b;
      EB_TRY(pin_acquire(c, b));
      const void* d = src(o, m);
      if (!d) return EB200_E_CUDA;
      EB_CUDA(c, cudaMemcpyAsync(c->pin[b], d, m, cudaMemcpyDeviceToHost, c->stream));
      EB_CUDA(c, cudaEventRecord(c->pin_ev[b], c->stream));
      c->pin_busy[b] = true;
      slot[issued % NP] = b;
      ++issued;
    }
    const int b = slot[i % NP];
    const size_t o = i * chunk, m = std::min(chunk, bytes - o);
    EB_CUDA(c, cudaEventSynchronize(c->pin_ev[b]));
    c->pin_busy[b] = false;
    parallel_memcpy((char*)host + o, c->pin[b], m, c->host_threads);
  }
  return 0;
}

}  // namespace eb
