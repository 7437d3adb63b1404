// C ABI of the shared-memory object store (csrc/store/store.cc).
// Role in the reference: Ray's plasma store, used via ray.put/ray.get of Arrow tables
// (reference python/raydp/spark/dataset.py:78-84, :222, :307-311).
#pragma once
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

// Create `/dev/shm/<name>` sized for meta_size + data_size bytes; returns 0 or -errno.
int rdp_store_create(const char* name, uint64_t data_size, uint64_t meta_size, void** handle_out);
// Map an existing sealed object; -EAGAIN if the producer has not sealed it yet.
int rdp_store_open(const char* name, int writable, void** handle_out);
int rdp_store_seal(void* handle);
void* rdp_store_data(void* handle);
void* rdp_store_meta(void* handle);
uint64_t rdp_store_data_size(void* handle);
uint64_t rdp_store_meta_size(void* handle);
uint64_t rdp_store_created_ns(void* handle);
int rdp_store_shrink(void* handle, uint64_t new_data_size);
int64_t rdp_store_incref(void* handle, int64_t delta);
// cudaHostRegister the mapping in this process (0, -ENOSYS without CUDA, -(1000+cudaError)).
int rdp_store_pin(void* handle);
int rdp_store_close(void* handle);
int rdp_store_unlink(const char* name);
int rdp_store_exists(const char* name);
int64_t rdp_store_unlink_prefix(const char* prefix);
int64_t rdp_store_usage(const char* prefix, uint64_t* bytes_out);
void rdp_store_parallel_copy(void* dst, const void* src, uint64_t n, int threads);

#ifdef __cplusplus
}
#endif
