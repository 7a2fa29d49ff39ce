/* C ABI of libtfdl_runtime.so - host-side runtime (no CUDA dependency).
 *
 * Return convention: 0 = ok; negative = failure, tfdl_rt_last_error() has the message
 * (-2: a peer failed / connection lost, -3: timeout). Handles are opaque pointers.
 * "ref:" notes name the line of /root/reference each piece stands in for.
 */
#ifndef TFDL_RUNTIME_H_
#define TFDL_RUNTIME_H_
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

const char* tfdl_rt_last_error(void);
int tfdl_rt_abi_version(void);

/* ---- crc32c (checkpoint + event-file checksums) ------------------------------------------ */
uint32_t tfdl_rt_crc32c(uint32_t crc, const void* data, uint64_t n);
uint32_t tfdl_rt_crc32c_mask(uint32_t crc);

/* ---- input pipeline helpers (ref: tf_dist_example.py:22-33 map/shuffle/batch) ------------ */
int tfdl_rt_shuffle_indices(const int64_t* src, int64_t n, int64_t buffer, uint64_t seed, int64_t* out);
/* the same shuffle, produced on demand: next() writes up to `max` more indices and returns the count (0 = done) */
void* tfdl_rt_shuffler_create(const int64_t* src, int64_t n, int64_t buffer, uint64_t seed);
int64_t tfdl_rt_shuffler_next(void* shuffler, int64_t* out, int64_t max);
void tfdl_rt_shuffler_destroy(void* shuffler);
int tfdl_rt_gather_rows(const void* src, int64_t num_rows, int64_t row_bytes, const int64_t* idx, int64_t n,
                        void* dst, int threads);
int tfdl_rt_gather_rows_u8_to_f32(const uint8_t* src, int64_t num_rows, int64_t row_elems, const int64_t* idx,
                                  int64_t n, float scale, float* dst, int threads);

/* k records [x_i | int32 labels_i] into a (pinned) ring; labels given as int32 (label_bytes 4) or int64 (8) */
int tfdl_rt_pack_records(const void* const* xs, const void* const* ys, int64_t k, int64_t x_bytes, int64_t n_labels,
                         int label_bytes, void* ring, int64_t rec_bytes, int threads);

/* k records gathered straight from dataset arrays: record r = [x rows idx[r*B .. r*B+B) | their labels as int32]
 * (labels stored as uint8, int32 or int64: label_bytes 1, 4, 8). shuffle + batch + staging in one pass. */
int tfdl_rt_gather_records(const void* x_src, int64_t num_rows, int64_t x_row_bytes, const void* y_src, int label_bytes,
                           const int64_t* idx, int64_t k, int64_t B, void* ring, int64_t rec_bytes, int threads);
/* the same gather, asynchronous (the prefetch of this input path): begin() posts the job to the worker pool and returns;
 * ready() = number of leading records complete; wait(n) returns once records [0, n) are complete (the caller copies
 * unclaimed parts itself instead of sleeping); end() completes and frees the job. x_src, y_src and ring must stay alive
 * until end(); idx is copied. */
void* tfdl_rt_gather_records_begin(const void* x_src, int64_t num_rows, int64_t x_row_bytes, const void* y_src,
                                   int label_bytes, const int64_t* idx, int64_t k, int64_t B, void* ring, int64_t rec_bytes);
int64_t tfdl_rt_gather_records_ready(void* job);
int tfdl_rt_gather_records_wait(void* job, int64_t n);
void tfdl_rt_gather_records_end(void* job);

/* ---- TCP control plane + RING collective (ref: README.md:21-23,65-68) -------------------- */
/* Binds the port of addrs[rank], connects ring/control/heartbeat channels, returns once every
 * peer has joined (or NULL after timeout_ms). addrs = training tasks in rank order. */
void* tfdl_rt_group_create(int rank, int world, const char* const* addrs, int timeout_ms, int hb_interval_ms,
                           int hb_timeout_ms);
void tfdl_rt_group_destroy(void* group);
int tfdl_rt_group_rank(void* group);
int tfdl_rt_group_world(void* group);
int tfdl_rt_group_failed_rank(void* group); /* -1 while everyone is alive */
/* word: a uint32 in pinned, device-mapped host memory that GPU kernels poll in their spin loops. The heartbeat thread
 * stores failed_rank + 1 into it as soon as a peer is declared dead (NULL detaches it before the memory is freed). */
void tfdl_rt_group_set_abort_word(void* group, uint32_t* word);
void tfdl_rt_group_set_timeout(void* group, int ms);
uint64_t tfdl_rt_group_bytes_sent(void* group);
int tfdl_rt_barrier(void* group);
int tfdl_rt_allgather(void* group, const void* in, uint64_t nbytes, void* out /* world*nbytes */);
int tfdl_rt_broadcast0(void* group, void* buf, uint64_t nbytes);
int tfdl_rt_ring_allreduce_f32(void* group, float* buf, int64_t n);
int tfdl_rt_ring_allreduce_f64(void* group, double* buf, int64_t n);

/* ---- TensorBundle checkpoints (ref: README.md:51 "chief saves checkpoints") --------------- */
void* tfdl_rt_bundle_writer_create(const char* prefix);
int tfdl_rt_bundle_writer_add(void* w, const char* key, int tf_dtype, int ndim, const int64_t* dims,
                              const void* data, uint64_t nbytes);
int tfdl_rt_bundle_writer_add_string(void* w, const char* key, const void* data, uint64_t nbytes);
int tfdl_rt_bundle_writer_finish(void* w); /* writes <prefix>.index and .data-00000-of-00001 atomically; frees w */
void tfdl_rt_bundle_writer_abort(void* w);
void* tfdl_rt_bundle_reader_open(const char* prefix);
void tfdl_rt_bundle_reader_close(void* r);
int tfdl_rt_bundle_reader_num_entries(void* r);
int tfdl_rt_bundle_reader_entry(void* r, int i, char* key, int key_cap, int* tf_dtype, int* ndim, int64_t* dims8,
                                uint64_t* nbytes);
int tfdl_rt_bundle_reader_read(void* r, int i, void* dst, uint64_t cap, uint64_t* written);

/* ---- TensorBoard event files (ref: README.md:51 "chief ... TensorBoard") ------------------ */
void* tfdl_rt_events_open(const char* logdir, const char* suffix);
int tfdl_rt_events_scalar(void* w, const char* tag, double value, int64_t step);
int tfdl_rt_events_flush(void* w);
const char* tfdl_rt_events_path(void* w);
void tfdl_rt_events_close(void* w);

#ifdef __cplusplus
}
#endif
#endif /* TFDL_RUNTIME_H_ */
