/* C ABI of libb200rt.so -- host-side runtime (no CUDA dependency).
 *
 * 1. Collective runtime: the Horovod-core equivalent used by `parallel/` when tensors live in host
 *    memory (the reference's `mpiexec -n 2 python train.py ...` plumbing config, train.py:18) and as
 *    the control plane for dynamic, name-keyed collectives.  One background cycle thread per
 *    process, name-keyed request table, rank-0 coordinator that releases a tensor only when every
 *    rank has asked for it (with dtype/shape/op validation), tensor fusion into a shared-memory
 *    fusion buffer, reduce-scatter + all-gather on host memory, stall warnings, Chrome-trace
 *    timeline.  Replaces Horovod's C++ core + MPI.
 * 2. HDF5 subset reader/writer for Keras `.h5` checkpoints (no libhdf5 on the box; N9).
 * 3. Bounding-box IoU (the Cython compute_overlap, N6) and bbox COCOeval (pycocotools, N7).
 */
#ifndef B200_RT_H_
#define B200_RT_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- collective runtime ------------------------------------------------------------------ */
enum { B200RT_ALLREDUCE = 0, B200RT_BROADCAST = 1, B200RT_ALLGATHER = 2 };
enum { B200RT_F32 = 0, B200RT_F64 = 1, B200RT_I32 = 2, B200RT_I64 = 3, B200RT_U8 = 4 };

/* job_id names the shared-memory segment (all ranks of a job pass the same string).
 * cycle_ms <= 0 -> 5 ms (HOROVOD_CYCLE_TIME); fusion_bytes <= 0 -> 64 MiB (HOROVOD_FUSION_THRESHOLD). */
int b200rt_init(int rank, int size, int local_rank, int local_size, const char* job_id, double cycle_ms,
                long long fusion_bytes);
int b200rt_shutdown(void);
int b200rt_initialized(void);
int b200rt_rank(void);
int b200rt_size(void);
int b200rt_local_rank(void);
int b200rt_local_size(void);

/* Enqueue an asynchronous collective; returns a handle (> 0) or a negative error.
 * allreduce: out[i] = postscale * sum_r in_r[i] (count elements).
 * broadcast: out = root's in (count elements).
 * allgather: out = concat_r in_r; `count` may differ per rank, out must hold the total. */
long long b200rt_enqueue(int op, const char* name, const void* in, void* out, long long count, int dtype, int root,
                         double postscale);
int b200rt_poll(long long handle);             /* 1 done, 0 pending, < 0 failed */
int b200rt_wait(long long handle);             /* 0 ok, < 0 failed (message via b200rt_last_error) */
long long b200rt_result_count(long long handle); /* allgather: total elements; releases nothing */
void b200rt_release(long long handle);
int b200rt_barrier(void);
const char* b200rt_last_error(void);

int b200rt_timeline_start(const char* path);   /* rank 0 writes a Chrome trace (HOROVOD_TIMELINE) */
int b200rt_timeline_stop(void);
/* GPU-path phases (device-timed with CUDA events by the engine) written into the same trace, on this clock. */
unsigned long long b200rt_now_ns(void);
int b200rt_timeline_active(void);
int b200rt_timeline_event(const char* phase, const char* tensor, unsigned long long t0_ns, unsigned long long t1_ns);
/* statistics: cycles run, responses executed, tensors fused, bytes reduced, stall warnings */
void b200rt_stats(long long out[5]);

/* ---- IoU (compute_overlap.pyx) ------------------------------------------------------------ */
/* boxes [n,4], query [k,4] float64 -> out [n,k]; +1 pixel convention when plus_one != 0. */
void b200rt_compute_overlap(const double* boxes, long long n, const double* query, long long k, double* out,
                            int plus_one);

/* ---- bbox COCOeval -------------------------------------------------------------------------- */
/* gt: [ng,7] = image_id, category_id, x, y, w, h, iscrowd(+2 * ignore);  gt_area [ng]
 * dt: [nd,7] = image_id, category_id, x, y, w, h, score
 * stats_out[12]: the 12 numbers pycocotools' summarize() prints for iouType='bbox'. */
int b200rt_cocoeval_bbox(const double* gt, const double* gt_area, long long ng, const double* dt, long long nd,
                         const long long* img_ids, long long n_img, const long long* cat_ids, long long n_cat,
                         double* stats_out);

/* ---- HDF5 subset ---------------------------------------------------------------------------- */
typedef struct B200H5File B200H5File;
B200H5File* b200h5_open(const char* path);                 /* read */
B200H5File* b200h5_create(const char* path);               /* write (flushed by close) */
int b200h5_close(B200H5File* f);
const char* b200h5_last_error(void);
/* Reading.  Paths are absolute ("/model_weights/conv1/conv1/kernel:0").  List calls return a
 * '\n'-separated string valid until the next call on the same file. */
const char* b200h5_list(B200H5File* f, const char* group);           /* child names */
int b200h5_is_group(B200H5File* f, const char* path);
const char* b200h5_attr_names(B200H5File* f, const char* path);
/* attribute / dataset info: class 0 = float, 1 = int, 2 = fixed string, 3 = vlen string */
int b200h5_attr_info(B200H5File* f, const char* path, const char* name, int* cls, int* elem_size, int* rank,
                     long long dims[8]);
long long b200h5_attr_read(B200H5File* f, const char* path, const char* name, void* buf, long long buf_bytes);
int b200h5_dataset_info(B200H5File* f, const char* path, int* cls, int* elem_size, int* rank, long long dims[8]);
long long b200h5_dataset_read(B200H5File* f, const char* path, void* buf, long long buf_bytes);
/* Writing. */
int b200h5_make_group(B200H5File* f, const char* path);
int b200h5_write_dataset(B200H5File* f, const char* path, int cls, int elem_size, int rank, const long long* dims,
                         const void* data);
int b200h5_write_attr(B200H5File* f, const char* path, const char* name, int cls, int elem_size, int rank,
                      const long long* dims, const void* data);

#ifdef __cplusplus
}
#endif
#endif /* B200_RT_H_ */
