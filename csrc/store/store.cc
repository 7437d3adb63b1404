// Shared-memory object store for raydp_b200 (S1 in SURVEY.md section 8).
//
// Replaces what the reference borrows from Ray's plasma store (reference
// python/raydp/spark/dataset.py:78-84,222 puts pyarrow Tables into plasma): every
// object is one POSIX shared-memory segment `/dev/shm/<name>` = [Header | meta | data].
// Producers create + fill + seal; consumers map it read-only and build zero-copy Arrow
// buffers over `data`.  Per-object segments (instead of one arena guarded by a shared
// mutex) mean a crashed producer can never leave a lock held, and deleting a "node" is
// unlinking a name prefix -- which is how the fault-injection tests lose objects.
//
// The data area is 256-byte aligned so a consumer can cudaHostRegister() the mapping and
// DMA Arrow column buffers to the GPU in place (ingestion kernel K5).
//
// C ABI only (loaded with ctypes); declarations live in include/raydp_b200_store.h.
#include <atomic>
#include <cerrno>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include <dirent.h>
#include <dlfcn.h>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

#include "../../include/raydp_b200_store.h"

namespace {

constexpr uint64_t kMagic = 0x5244504232303053ull;  // "RDPB200S"
constexpr uint64_t kAlign = 256;

struct Header {
  uint64_t magic;
  uint32_t version;
  std::atomic<uint32_t> sealed;
  std::atomic<int64_t> refcount;
  uint64_t meta_size;
  uint64_t data_size;
  uint64_t data_offset;
  uint64_t total_size;
  uint64_t created_ns;
  uint64_t pinned;  // bookkeeping only; registration is per process
  char reserved[kAlign - 72];
};
static_assert(sizeof(Header) == kAlign, "header must be one alignment unit");

struct Mapping {
  void* base;
  uint64_t total;
  bool writable;
  bool pinned;
};

inline uint64_t round_up(uint64_t v, uint64_t a) { return (v + a - 1) / a * a; }

uint64_t now_ns() {
  timespec ts;
  clock_gettime(CLOCK_REALTIME, &ts);
  return uint64_t(ts.tv_sec) * 1000000000ull + uint64_t(ts.tv_nsec);
}

std::string shm_path(const char* name) {
  std::string n(name);
  if (n.empty() || n[0] != '/') n = "/" + n;
  return n;
}

// cudaHostRegister / cudaHostUnregister resolved lazily so the store builds and runs on
// hosts without CUDA.
typedef int (*host_register_fn)(void*, size_t, unsigned int);
typedef int (*host_unregister_fn)(void*);
host_register_fn g_register = nullptr;
host_unregister_fn g_unregister = nullptr;
std::atomic<int> g_cuda_state{0};  // 0 unknown, 1 ok, -1 unavailable

bool load_cuda() {
  int st = g_cuda_state.load();
  if (st != 0) return st > 0;
  const char* names[] = {"libcudart.so.12", "libcudart.so", "libcudart.so.13", nullptr};
  void* h = nullptr;
  for (int i = 0; names[i] && !h; ++i) h = dlopen(names[i], RTLD_NOW | RTLD_GLOBAL);
  if (h) {
    g_register = reinterpret_cast<host_register_fn>(dlsym(h, "cudaHostRegister"));
    g_unregister = reinterpret_cast<host_unregister_fn>(dlsym(h, "cudaHostUnregister"));
  }
  bool ok = g_register && g_unregister;
  g_cuda_state.store(ok ? 1 : -1);
  return ok;
}

}  // namespace

extern "C" {

int rdp_store_create(const char* name, uint64_t data_size, uint64_t meta_size, void** handle_out) {
  if (!name || !handle_out) return -EINVAL;
  std::string p = shm_path(name);
  int fd = shm_open(p.c_str(), O_CREAT | O_EXCL | O_RDWR, 0600);
  if (fd < 0) return -errno;
  uint64_t data_off = round_up(sizeof(Header) + meta_size, kAlign);
  uint64_t total = round_up(data_off + (data_size ? data_size : 1), 4096);
  if (ftruncate(fd, off_t(total)) != 0) {
    int e = errno;
    close(fd);
    shm_unlink(p.c_str());
    return -e;
  }
  void* base = mmap(nullptr, total, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
  close(fd);
  if (base == MAP_FAILED) {
    int e = errno;
    shm_unlink(p.c_str());
    return -e;
  }
  Header* h = new (base) Header();
  h->magic = kMagic;
  h->version = 1;
  h->sealed.store(0);
  h->refcount.store(1);
  h->meta_size = meta_size;
  h->data_size = data_size;
  h->data_offset = data_off;
  h->total_size = total;
  h->created_ns = now_ns();
  h->pinned = 0;
  *handle_out = new Mapping{base, total, true, false};
  return 0;
}

int rdp_store_open(const char* name, int writable, void** handle_out) {
  if (!name || !handle_out) return -EINVAL;
  std::string p = shm_path(name);
  int fd = shm_open(p.c_str(), writable ? O_RDWR : O_RDONLY, 0600);
  if (fd < 0) return -errno;
  struct stat st;
  if (fstat(fd, &st) != 0 || uint64_t(st.st_size) < sizeof(Header)) {
    int e = errno ? errno : EIO;
    close(fd);
    return -e;
  }
  // PROT_WRITE mappings are needed for cudaHostRegister on some drivers and for refcounts.
  int prot = writable ? (PROT_READ | PROT_WRITE) : PROT_READ;
  void* base = mmap(nullptr, size_t(st.st_size), prot, MAP_SHARED, fd, 0);
  close(fd);
  if (base == MAP_FAILED) return -errno;
  Header* h = static_cast<Header*>(base);
  if (h->magic != kMagic || h->total_size != uint64_t(st.st_size)) {
    munmap(base, size_t(st.st_size));
    return -EBADMSG;
  }
  if (!h->sealed.load(std::memory_order_acquire)) {
    munmap(base, size_t(st.st_size));
    return -EAGAIN;  // producer has not sealed it yet
  }
  *handle_out = new Mapping{base, uint64_t(st.st_size), writable != 0, false};
  return 0;
}

int rdp_store_seal(void* handle) {
  Mapping* m = static_cast<Mapping*>(handle);
  if (!m || !m->writable) return -EINVAL;
  static_cast<Header*>(m->base)->sealed.store(1, std::memory_order_release);
  return 0;
}

void* rdp_store_data(void* handle) {
  Mapping* m = static_cast<Mapping*>(handle);
  Header* h = static_cast<Header*>(m->base);
  return static_cast<char*>(m->base) + h->data_offset;
}

void* rdp_store_meta(void* handle) {
  Mapping* m = static_cast<Mapping*>(handle);
  return static_cast<char*>(m->base) + sizeof(Header);
}

uint64_t rdp_store_data_size(void* handle) {
  return static_cast<Header*>(static_cast<Mapping*>(handle)->base)->data_size;
}

uint64_t rdp_store_meta_size(void* handle) {
  return static_cast<Header*>(static_cast<Mapping*>(handle)->base)->meta_size;
}

uint64_t rdp_store_created_ns(void* handle) {
  return static_cast<Header*>(static_cast<Mapping*>(handle)->base)->created_ns;
}

int rdp_store_shrink(void* handle, uint64_t new_data_size) {
  // Producers that over-estimated (Arrow IPC framing) record the real payload size.
  Mapping* m = static_cast<Mapping*>(handle);
  if (!m || !m->writable) return -EINVAL;
  Header* h = static_cast<Header*>(m->base);
  if (new_data_size > h->data_size) return -ERANGE;
  h->data_size = new_data_size;
  return 0;
}

int64_t rdp_store_incref(void* handle, int64_t delta) {
  Mapping* m = static_cast<Mapping*>(handle);
  if (!m || !m->writable) return -EINVAL;
  return static_cast<Header*>(m->base)->refcount.fetch_add(delta) + delta;
}

int rdp_store_pin(void* handle) {
  // Page-lock the whole mapping for this process so cudaMemcpyAsync from Arrow buffers
  // inside it is a true async DMA.  Returns 0 on success, -ENOSYS without CUDA.
  Mapping* m = static_cast<Mapping*>(handle);
  if (!m) return -EINVAL;
  if (m->pinned) return 0;
  if (!load_cuda()) return -ENOSYS;
  unsigned flags = 0x01 /*portable*/;
  if (!m->writable) flags |= 0x08 /*cudaHostRegisterReadOnly*/;
  int rc = g_register(m->base, size_t(m->total), flags);
  if (rc != 0 && !m->writable) rc = g_register(m->base, size_t(m->total), 0x01);
  if (rc != 0) return -(1000 + rc);
  m->pinned = true;
  return 0;
}

int rdp_store_close(void* handle) {
  Mapping* m = static_cast<Mapping*>(handle);
  if (!m) return -EINVAL;
  if (m->pinned && g_unregister) g_unregister(m->base);
  munmap(m->base, size_t(m->total));
  delete m;
  return 0;
}

int rdp_store_unlink(const char* name) {
  std::string p = shm_path(name);
  return shm_unlink(p.c_str()) == 0 ? 0 : -errno;
}

int rdp_store_exists(const char* name) {
  std::string p = std::string("/dev/shm") + shm_path(name);
  struct stat st;
  return stat(p.c_str(), &st) == 0 ? 1 : 0;
}

int64_t rdp_store_unlink_prefix(const char* prefix) {
  // Drops every object whose name starts with `prefix`: session teardown and the
  // "node died" path of the fault-injection tests.
  DIR* d = opendir("/dev/shm");
  if (!d) return -errno;
  std::string pre(prefix);
  if (!pre.empty() && pre[0] == '/') pre = pre.substr(1);
  int64_t n = 0;
  std::vector<std::string> victims;
  while (dirent* e = readdir(d)) {
    if (strncmp(e->d_name, pre.c_str(), pre.size()) == 0) victims.emplace_back(e->d_name);
  }
  closedir(d);
  for (auto& v : victims) {
    if (shm_unlink(("/" + v).c_str()) == 0) ++n;
  }
  return n;
}

int64_t rdp_store_usage(const char* prefix, uint64_t* bytes_out) {
  DIR* d = opendir("/dev/shm");
  if (!d) return -errno;
  std::string pre(prefix);
  if (!pre.empty() && pre[0] == '/') pre = pre.substr(1);
  int64_t n = 0;
  uint64_t bytes = 0;
  while (dirent* e = readdir(d)) {
    if (strncmp(e->d_name, pre.c_str(), pre.size()) != 0) continue;
    struct stat st;
    std::string p = std::string("/dev/shm/") + e->d_name;
    if (stat(p.c_str(), &st) == 0) {
      bytes += uint64_t(st.st_size);
      ++n;
    }
  }
  closedir(d);
  if (bytes_out) *bytes_out = bytes;
  return n;
}

void rdp_store_parallel_copy(void* dst, const void* src, uint64_t n, int threads) {
  // memcpy split over threads: one core cannot saturate DDR5 when filling a block.
  if (threads <= 1 || n < (8ull << 20)) {
    memcpy(dst, src, n);
    return;
  }
  uint64_t chunk = round_up((n + threads - 1) / threads, 4096);
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; ++t) {
    uint64_t off = uint64_t(t) * chunk;
    if (off >= n) break;
    uint64_t len = (off + chunk > n) ? n - off : chunk;
    pool.emplace_back([=] { memcpy(static_cast<char*>(dst) + off, static_cast<const char*>(src) + off, len); });
  }
  for (auto& th : pool) th.join();
}

}  // extern "C"
