// Probe of the host <-> device paths a pageable-memory caller can take (round 3).
//   nvcc -O2 -o /tmp/probe scripts/r3/probe_host_link.cu && /tmp/probe
#include <cuda_runtime.h>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>
static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
static void par_memcpy(char *d, const char *s, size_t n, int nt) {
  if (nt <= 1) { memcpy(d, s, n); return; }
  std::vector<std::thread> th;
  size_t per = (n / nt + 4095) & ~(size_t)4095;
  for (int t = 0; t < nt; ++t) {
    size_t o = (size_t)t * per;
    if (o >= n) break;
    size_t len = std::min(per, n - o);
    th.emplace_back([=] { memcpy(d + o, s + o, len); });
  }
  for (auto &t : th) t.join();
}
int main() {
  const size_t N = (size_t)512 << 20;
  char *dev; cudaMalloc(&dev, N);
  char *pin; cudaHostAlloc(&pin, N, cudaHostAllocDefault);
  memset(pin, 1, N);
  char *pg = (char *)malloc(N); memset(pg, 2, N);
  cudaStream_t s; cudaStreamCreate(&s);
  auto rate = [&](const char *name, auto fn) { fn(); double t0 = now(); fn(); double dt = now() - t0; printf("%-46s %7.2f GB/s (%.1f ms)\n", name, N / dt / 1e9, dt * 1e3); };
  rate("pinned H2D", [&] { cudaMemcpyAsync(dev, pin, N, cudaMemcpyHostToDevice, s); cudaStreamSynchronize(s); });
  rate("pinned D2H", [&] { cudaMemcpyAsync(pin, dev, N, cudaMemcpyDeviceToHost, s); cudaStreamSynchronize(s); });
  rate("pageable H2D (driver staging)", [&] { cudaMemcpyAsync(dev, pg, N, cudaMemcpyHostToDevice, s); cudaStreamSynchronize(s); });
  rate("pageable D2H, touched pages", [&] { cudaMemcpyAsync(pg, dev, N, cudaMemcpyDeviceToHost, s); cudaStreamSynchronize(s); });
  {
    char *fresh = (char *)malloc(N);
    double t0 = now(); cudaMemcpyAsync(fresh, dev, N, cudaMemcpyDeviceToHost, s); cudaStreamSynchronize(s); double dt = now() - t0;
    printf("%-46s %7.2f GB/s (%.1f ms)\n", "pageable D2H, untouched pages (one shot)", N / dt / 1e9, dt * 1e3);
    free(fresh);
  }
  for (int nt : {1, 2, 4, 8}) {
    char name[64]; snprintf(name, sizeof name, "memcpy pageable -> pinned, %d thread(s)", nt);
    rate(name, [&] { par_memcpy(pin, pg, N, nt); });
  }
  for (int nt : {1, 4}) {
    char *fresh = (char *)malloc(N);
    double t0 = now(); par_memcpy(fresh, pin, N, nt); double dt = now() - t0;
    printf("memcpy pinned -> untouched pageable, %d thr      %7.2f GB/s (%.1f ms)\n", nt, N / dt / 1e9, dt * 1e3);
    free(fresh);
  }
  {
    double t0 = now(); cudaError_t e = cudaHostRegister(pg, N, cudaHostRegisterDefault); double dt = now() - t0;
    printf("cudaHostRegister of 512 MiB: %s, %.1f ms (%.2f GB/s)\n", cudaGetErrorString(e), dt * 1e3, N / dt / 1e9);
    if (e == cudaSuccess) {
      rate("registered H2D", [&] { cudaMemcpyAsync(dev, pg, N, cudaMemcpyHostToDevice, s); cudaStreamSynchronize(s); });
      t0 = now(); cudaHostUnregister(pg); printf("cudaHostUnregister: %.1f ms\n", (now() - t0) * 1e3);
    }
  }
  // staged pipeline: chunks through 4 pinned buffers, nt copy threads
  for (int nt : {1, 4}) for (size_t chunk : {(size_t)4 << 20, (size_t)16 << 20}) {
    const int NB = 4; cudaEvent_t ev[NB]; for (auto &e : ev) cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
    auto h2d = [&] { size_t off = 0; int i = 0; while (off < N) { size_t len = std::min(chunk, N - off); char *b = pin + (size_t)(i % NB) * chunk; if (i >= NB) cudaEventSynchronize(ev[i % NB]); par_memcpy(b, pg + off, len, nt); cudaMemcpyAsync(dev + off, b, len, cudaMemcpyHostToDevice, s); cudaEventRecord(ev[i % NB], s); off += len; ++i; } cudaStreamSynchronize(s); };
    char name[80]; snprintf(name, sizeof name, "staged H2D, %zu MiB chunks, %d thread(s)", chunk >> 20, nt);
    rate(name, h2d);
    auto d2h = [&] { size_t nchunks = (N + chunk - 1) / chunk; for (size_t i = 0; i < nchunks + NB; ++i) { if (i >= (size_t)NB) { size_t j = i - NB; cudaEventSynchronize(ev[j % NB]); size_t off = j * chunk; par_memcpy(pg + off, pin + (j % NB) * chunk, std::min(chunk, N - off), nt); } if (i < nchunks) { size_t off = i * chunk; cudaMemcpyAsync(pin + (i % NB) * chunk, dev + off, std::min(chunk, N - off), cudaMemcpyDeviceToHost, s); cudaEventRecord(ev[i % NB], s); } } };
    snprintf(name, sizeof name, "staged D2H, %zu MiB chunks, %d thread(s)", chunk >> 20, nt);
    rate(name, d2h);
  }
  printf("host threads: %u\n", std::thread::hardware_concurrency());
  return 0;
}
