// Test shim: exposes voxel_game_b200/csrc/host_queues.h (pure host C++, no CUDA) to ctypes.
#include "../voxel_game_b200/csrc/host_queues.h"
using namespace vxb;
extern "C" {
void* qm_new() { return new QueueMapForced(); }
void qm_del(void* q) { delete (QueueMapForced*)q; }
int qm_push(void* q, const uint32_t* k, int v) { return ((QueueMapForced*)q)->push(Key3{{k[0], k[1], k[2]}}, v != 0); }
int qm_pop(void* q, uint32_t* k, int* v) { Key3 key; bool val; if (!((QueueMapForced*)q)->pop(key, val)) return 0; k[0]=key.v[0];k[1]=key.v[1];k[2]=key.v[2]; *v = val; return 1; }
uint64_t qm_len(void* q) { return ((QueueMapForced*)q)->len(); }
void* qs_new() { return new QueueSet4(); }
void qs_del(void* q) { delete (QueueSet4*)q; }
void qs_push(void* q, const uint32_t* k) { ((QueueSet4*)q)->push(Key4{{k[0], k[1], k[2], k[3]}}); }
int qs_pop(void* q, uint32_t* k) { Key4 key; if (!((QueueSet4*)q)->pop(key)) return 0; for (int i=0;i<4;i++) k[i]=key.v[i]; return 1; }
uint64_t qs_len(void* q) { return ((QueueSet4*)q)->len(); }
void* uq_new() { return new VoxelUpdateQueue(); }
void uq_del(void* q) { delete (VoxelUpdateQueue*)q; }
void uq_push_all(void* q, const uint32_t* k) { ((VoxelUpdateQueue*)q)->push_all(Key3{{k[0], k[1], k[2]}}); }
void uq_push_prio(void* q, const uint32_t* k, int32_t p) { ((VoxelUpdateQueue*)q)->push_with_priority(Key3{{k[0], k[1], k[2]}}, p); }
void uq_swap(void* q) { ((VoxelUpdateQueue*)q)->swap_queue_set(); }
int uq_pop(void* q, uint32_t* k, int32_t* p) { Key3 key; if (!((VoxelUpdateQueue*)q)->pop(key, *p)) return 0; k[0]=key.v[0];k[1]=key.v[1];k[2]=key.v[2]; return 1; }
void uq_idx(void* q, uint32_t* out) { ((VoxelUpdateQueue*)q)->queue_set_idx(out); }
uint64_t uq_len(void* q) { return ((VoxelUpdateQueue*)q)->len(); }
int uq_empty(void* q) { return ((VoxelUpdateQueue*)q)->is_empty(); }
}
