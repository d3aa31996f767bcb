/* C ABI of libpdl_host.so -- the host-side runtime (no CUDA dependency).
 *
 * Rendezvous store, CPU ("gloo"-role) collectives and gradient-bucket assignment.  These are the
 * native equivalents of what the reference borrows from torch's C++ (TCPStore, ProcessGroupGloo,
 * compute_bucket_assignment_by_size); reference call sites are cited per function.
 * All functions are thread-compatible; handles are opaque pointers.
 */
#ifndef PDL_B200_HOST_H_
#define PDL_B200_HOST_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- key/value store (rank 0 serves; reference: init_process_group at mpspawn_dist.py:49-54 env://,
 *      example_mp.py:37-42 tcp://) ------------------------------------------------------------- */
void* pdl_store_server_start(const char* host, int port, int* out_port); /* port 0 = ephemeral; NULL on bind failure */
void pdl_store_server_stop(void* server);
void* pdl_store_client_connect(const char* host, int port, double timeout_s); /* retries until the server is up */
void pdl_store_client_close(void* client);
/* return codes: 0 ok, 1 timeout, -1 connection lost */
int pdl_store_set(void* client, const char* key, const void* val, uint64_t n);
int pdl_store_get(void* client, const char* key, double timeout_s, void** out, uint64_t* out_n); /* blocks; *out is malloc'ed */
int pdl_store_add(void* client, const char* key, int64_t delta, int64_t* result);
int pdl_store_wait(void* client, const char** keys, int nkeys, double timeout_s);
int pdl_store_check(void* client, const char** keys, int nkeys); /* 0 when every key exists */
int pdl_store_delete(void* client, const char* key);             /* 1 when a key was removed */
int64_t pdl_store_num_keys(void* client);
int pdl_store_compare_set(void* client, const char* key, const void* expect, uint64_t en, const void* desired, uint64_t dn, void** out,
                          uint64_t* out_n);
void pdl_free(void* p);

/* ---- CPU collectives over a TCP mesh (reference: backend names in README.md:133; BASELINE config 1) ---- */
/* dtype: 0 f32, 1 f64, 2 i32, 3 i64, 4 u8.  op: 0 sum, 1 prod, 2 min, 3 max.  ranks are group-local. */
void* pdl_cpu_comm_create(void* store_client, const char* prefix, int rank, int world, const char* my_host, double timeout_s);
void pdl_cpu_comm_destroy(void* comm);
int pdl_cpu_allreduce(void* comm, void* buf, uint64_t count, int dtype, int op); /* ring reduce-scatter + allgather: identical bytes on every rank */
int pdl_cpu_broadcast(void* comm, void* buf, uint64_t nbytes, int root);
int pdl_cpu_allgather(void* comm, const void* send, void* recv, uint64_t nbytes);
int pdl_cpu_barrier(void* comm);
int pdl_cpu_send(void* comm, const void* buf, uint64_t nbytes, int dst);
int pdl_cpu_recv(void* comm, void* buf, uint64_t nbytes, int src);

/* ---- bucket assignment (reference: DDP constructor at launch_dist.py:61; torch's
 *      compute_bucket_assignment_by_size contract) ------------------------------------------------ */
int pdl_bucket_assign(const int64_t* sizes, const int32_t* kinds, int n, const int64_t* limits, int nlimits, int32_t* bucket_of);

/* ---- reducer core (reference: torch's C++ Reducer behind DistributedDataParallel, launch_dist.py:61): which buckets
 *      may be launched when a gradient becomes ready (all members ready AND every earlier bucket launched), what is
 *      left at the end of backward, the arrival order of the first iteration.  Tensors and kernels stay with the caller.
 *      mark_ready returns the number of buckets written to `out`, -1 unknown parameter, -2 second gradient for a
 *      parameter in one iteration, -3 `out` too small. ------------------------------------------------------------- */
void* pdl_reducer_create(int num_params, const int32_t* bucket_of, int num_buckets);
void pdl_reducer_destroy(void* reducer);
int pdl_reducer_set_layout(void* reducer, int num_params, const int32_t* bucket_of, int num_buckets);
void pdl_reducer_prepare(void* reducer, int record_arrival);
int pdl_reducer_mark_ready(void* reducer, int param, int32_t* out, int cap);
int pdl_reducer_finalize(void* reducer, int32_t* out, int cap);
int pdl_reducer_is_ready(void* reducer, int param);
int pdl_reducer_next_bucket(void* reducer);
int pdl_reducer_arrival(void* reducer, int32_t* out, int cap);

#ifdef __cplusplus
}
#endif
#endif /* PDL_B200_HOST_H_ */
