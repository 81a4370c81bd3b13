/*
 * gmt_b200.h -- C ABI of libgmt_b200.so, the sm_100a native layer of the
 * B200 distributed-array / halo-exchange runtime.
 *
 * Conventions
 *   - every entry point returns 0 on success, a positive cudaError_t value when
 *     a CUDA call failed, or a negative GMT_E* code for argument errors;
 *     gmt_last_error() returns a thread-local message for the last failure.
 *   - pointers are device pointers unless the name says host; `stream` is a
 *     cudaStream_t passed as void* (NULL = legacy default stream).
 *   - all 2-D arrays are COLUMN-MAJOR (first index fastest), the layout the
 *     reference uses (mpi_stencil2d_sycl.cc:40-43): element (i,j) of an array
 *     with leading dimension ld lives at i + j*ld.
 *   - nothing here calls MPI or NCCL.  Cross-GPU traffic is plain loads/stores
 *     (or multimem.*) on peer-mapped addresses supplied by the caller.
 */
#ifndef GMT_B200_H
#define GMT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* gmt_stream_t;

enum {
  GMT_OK = 0,
  GMT_EINVAL = -1,    /* bad argument (size, alignment, enum) */
  GMT_ETOOBIG = -2,   /* size beyond a documented maximum */
  GMT_EUNSUPPORTED = -3,
  GMT_ETIMEOUT = -4   /* an in-kernel bounded spin expired (peer never signalled) */
};

enum { GMT_SPACE_DEVICE = 0, GMT_SPACE_MANAGED = 1, GMT_SPACE_PINNED = 2 };

/* ------------------------------------------------------------------ errors */
const char* gmt_last_error(void);
int gmt_version(void);

/* ------------------------------------------- runtime / device introspection
 * Replaces set_rank_device (mpi_daxpy.cc:36-62), device_get_vendor_id
 * (mpi_stencil_gt.cc:151-152), cudaMemGetInfo report (mpi_daxpy_nvtx.cc:201-205)
 * and the PTRINFO / MEMINFO helpers (cuda_error.h:66-136).                  */
typedef struct {
  char name[256];
  uint64_t total_global_mem;
  int pci_domain, pci_bus, pci_device;
  int sm_count, cc_major, cc_minor;
  int clock_khz, mem_clock_khz, l2_bytes;
  int managed_memory, concurrent_managed_access;
  int max_smem_optin;
  unsigned char uuid[16];
} gmt_device_props;

int gmt_device_count(int* count);
int gmt_set_device(int dev);
int gmt_get_device(int* dev);
int gmt_device_props_get(int dev, gmt_device_props* out);
int gmt_mem_info(uint64_t* free_bytes, uint64_t* total_bytes);
/* type: 0 unregistered, 1 host, 2 device, 3 managed, -1 NULL (cuda_error.h:66-96) */
int gmt_ptr_info(const void* p, int* type, int* device);
/* loc: -3 not a CUDA pointer, -2 unmanaged, -1 CPU, >=0 device id, -4 invalid
 * (cuda_error.h:99-136) */
int gmt_mem_preferred_location(const void* p, size_t bytes, int* loc);
int gmt_can_access_peer(int dev, int peer, int* ok);
int gmt_enable_peer_access(int peer);
int gmt_device_synchronize(void);
int gmt_stream_synchronize(gmt_stream_t stream);

/* ------------------------------------------------------------ memory spaces
 * device / managed / pinned-host allocation (mpi_daxpy_nvtx.cc:177-199) and
 * CUDA-IPC export so peers can map a device allocation.                     */
int gmt_alloc(int space, size_t bytes, void** out);
int gmt_free(int space, void* p);
/* kind: 0 H2H, 1 H2D, 2 D2H, 3 D2D, 4 default */
int gmt_memcpy_async(void* dst, const void* src, size_t bytes, int kind, gmt_stream_t stream);
int gmt_memset_async(void* dst, int value, size_t bytes, gmt_stream_t stream);
int gmt_ipc_get_handle(void* dev_ptr, unsigned char handle[64]);
int gmt_ipc_open_handle(const unsigned char handle[64], void** out);
int gmt_ipc_close_handle(void* mapped);

/* --------------------------------------------------- tracing (daxpy_nvtx.cu:65-105,
 * mpi_daxpy_nvtx.cc:167-328): NVTX push/pop ranges + profiler capture window */
int gmt_nvtx_push(const char* name);
int gmt_nvtx_pop(void);
int gmt_profiler_start(void);
int gmt_profiler_stop(void);

/* ---------------------------------------------------------- local fp64 ops */
/* bytes of zero-initialised scratch every reduction-type entry point needs;
 * the same scratch may be reused by back-to-back calls on ONE stream.       */
size_t gmt_reduce_scratch_bytes(void);

/* y <- a*x + y   (replaces cublasDaxpy, mpi_daxpy_nvtx.cc:245) */
int gmt_daxpy_f64(int64_t n, double a, const double* x, double* y, gmt_stream_t stream);
/* y <- a*x + y and *out_sum <- sum(y_new), one pass over memory
 * (replaces daxpy + D2H + host loop, mpi_daxpy_nvtx.cc:245-268) */
int gmt_daxpy_sum_f64(int64_t n, double a, const double* x, double* y, double* out_sum,
                      void* scratch, gmt_stream_t stream);
/* Same result as gmt_daxpy_sum_f64 for callers that cannot allocate device memory.
 * (No counterpart in the reference.)  Meant for programs whose own arrays or memory pool
 * already fill the GPU, so a late cudaMalloc can fail, and for time-stepping loops where an
 * allocator call -- which may synchronise the whole device -- is not acceptable.  The scratch
 * is a fixed array inside the library image (gmt_noalloc_scratch_bytes() = 64.25 KiB, resident
 * on every GPU that loads the library until the process exits); the caller passes inputs and
 * outputs only, and nothing in the call allocates or frees.
 *   - n is limited to gmt_noalloc_max_n() = 33,554,432 elements (one partial per
 *     4096-element chunk must fit the array); larger n returns GMT_ETOOBIG.
 *   - The first call on each device looks up the array's device address; make that first
 *     call before any stream capture (a first call on a capturing stream returns
 *     GMT_EUNSUPPORTED).  The address is kept for the life of the process: do not call
 *     cudaDeviceReset() between calls.
 *   - All calls on a device share the array, which keeps the previous call's contents; the
 *     counter that must start at zero is cleared on `stream` by every call.  Calls that overlap
 *     on the GPU (unordered streams, concurrent host threads) race on it: order them.
 *   - The partials are added in a fixed order (strided over the threads of the last block, then
 *     a shuffle tree): the sum is bit-reproducible run to run, though not a left-to-right sum. */
int gmt_daxpy_sum_noalloc_f64(int64_t n, double a, const double* x, double* y, double* out_sum,
                              gmt_stream_t stream);
int64_t gmt_noalloc_max_n(void);
size_t gmt_noalloc_scratch_bytes(void);
/* *out <- sum(x)  (host loops mpi_daxpy_nvtx.cc:253-256,296-298) */
int gmt_sum_f64(int64_t n, const double* x, double* out, void* scratch, gmt_stream_t stream);
/* *out <- sum((a-b)^2)  (gt::sum_squares mpi_stencil_gt.cc:222; diff_norm
 * mpi_stencil2d_sycl.cc:165-181).  Caller takes the sqrt.                   */
int gmt_sumsq_diff_f64(int64_t n, const double* a, const double* b, double* out, void* scratch,
                       gmt_stream_t stream);
/* 2-D variant on strided column-major views: sum over (i<nx, j<ny) of (a(i,j)-b(i,j))^2 */
int gmt_sumsq_diff2d_f64(int64_t nx, int64_t ny, const double* a, int64_t lda, const double* b,
                         int64_t ldb, double* out, void* scratch, gmt_stream_t stream);
/* axis sum of a column-major (nx,ny,ld) array: axis 0 -> out[j]=sum_i z(i,j), j<n_out;
 * axis 1 -> out[i]=sum_j z(i,j), i<n_out  (gt::sum_axis_to mpi_stencil2d_gt.cc:611,620) */
int gmt_axis_sum_f64(int64_t nx, int64_t ny, int64_t ld, const double* z, int axis, int64_t n_out,
                     double* out, gmt_stream_t stream);
/* x[:] <- value  (fill ctor mpi_stencil2d_gt.cc:598) */
int gmt_fill_f64(int64_t n, double value, double* x, gmt_stream_t stream);
/* x[:] <- value for 64-bit words (epoch flags set from a stream) */
int gmt_fill_u64(int64_t n, uint64_t value, uint64_t* x, gmt_stream_t stream);
/* x[i] <- (start + i*step)*mul   -- daxpy initialisers (mpi_daxpy_nvtx.cc:209-212) */
int gmt_iota_f64(int64_t n, double start, double step, double mul, double* x, gmt_stream_t stream);
/* x[i-1] <- (rank*i)/N in 64-bit integer arithmetic, i = 1..n: the slot values of the in-place
 * gather test (mpigatherinplace.f90:34-37) */
int gmt_slot_pattern_f64(int64_t n, int64_t rank, int64_t N, double* x, gmt_stream_t stream);

/* 5-point 4th-order first derivative, out[i] = scale * sum_k c_k in[i+k], k=0..4
 * (stencil1d_5, mpi_stencil_gt.cc:54-59).  in has n+4 elements.
 * variant: 0 = one thread per point (mirrors a generic expression kernel),
 *          1 = 256-bit vectorised streaming kernel,
 *          2 = TMA: cp.async.bulk + mbarrier pipeline (1-D); for the 2-D dim-1 stencil,
 *              tensor-map loads + 2-CTA clusters, shared halo columns fetched once by TMA multicast
 *              (2-D dim 0 has no TMA variant and runs variant 1).             */
int gmt_stencil1d_f64(int64_t n, const double* in, double* out, double scale, int variant,
                      gmt_stream_t stream);
/* same stencil applied along dim 0 (contiguous) or dim 1 (strided) of a column-major
 * 2-D array.  dim 0: z is (nx+4, ny) with ldz >= nx+4; dim 1: z is (nx, ny+4) with
 * ldz >= nx.  dz is (nx, ny) with lddz >= nx.
 * (stencil2d_1d_5_d0/_d1, mpi_stencil2d_gt.cc:84-110)                        */
int gmt_stencil2d_f64(int dim, int64_t nx, int64_t ny, const double* z, int64_t ldz, double* dz,
                      int64_t lddz, double scale, int variant, gmt_stream_t stream);
/* variant 2 of the dim-1 stencil on its own: 2-D TMA tensor-map loads + thread-block clusters
 * of two CTAs; the 4 halo columns between their tiles are fetched once and delivered to both
 * CTAs' shared memory by TMA multicast (no reference counterpart; see csrc/stencil_tma.cu for the shape/alignment
 * requirements).  z must be a plain device allocation of the current GPU -- the TMA multicast loads read only this
 * GPU's memory; a peer / IPC / symmetric-memory mapping or managed memory returns GMT_EUNSUPPORTED.            */
int gmt_stencil2d_d1_cluster_f64(int64_t nx, int64_t ny, const double* z, int64_t ldz, double* out,
                                 int64_t ldo, double scale, gmt_stream_t stream);
/* cross stencil on a (nx+4, ny+4) array ghosted in both dims:
 * out(i,j) = sx * d/dx + sy * d/dy   (BASELINE config 3, no reference counterpart) */
int gmt_stencil2d_xy_f64(int64_t nx, int64_t ny, const double* z, int64_t ldz, double* out,
                         int64_t ldo, double sx, double sy, gmt_stream_t stream);

/* strided 2-D block copy dst(i,j) = src(i,j), i<rows (contiguous), j<cols: the pack /
 * unpack kernel of the baseline exchange (mpi_stencil2d_gt.cc:166,172,239,251;
 * buf_from_view / buf_to_view mpi_stencil2d_sycl.cc:82-116)                  */
int gmt_copy2d_f64(int64_t rows, int64_t cols, const double* src, int64_t ld_src, double* dst,
                   int64_t ld_dst, gmt_stream_t stream);

/* analytic initialisers, device side (host loops mpi_stencil_gt.cc:177-196,
 * mpi_stencil2d_gt.cc:445-497).
 * 1-D: y[g] = x^3 for every g in [0, n+2*nb) with x = x0 + (g-nb)*dx (so ghosts get the
 *      analytic continuation; the caller overwrites interior-facing ghosts by exchange),
 *      optional actual[i] = 3x^2.                                            */
int gmt_init_stencil1d_f64(int64_t n, int nb, double x0, double dx, double* y, double* actual,
                           int fill_left_ghost, int fill_right_ghost, gmt_stream_t stream);
/* 2-D: z(i,j) = x^3 + y^2 on the ghosted array (gx, gy ghost widths per dim), ghost cells
 *      filled analytically only on the sides whose bit is set in ghost_mask
 *      (1 = low x, 2 = high x, 4 = low y, 8 = high y), others set to `poison`.
 *      deriv_mode: 0 none, 1 actual = 3x^2, 2 actual = 2y, 3 actual = sx*3x^2 + sy*2y     */
int gmt_init_stencil2d_f64(int64_t nx, int64_t ny, int gx, int gy, double x0, double y0, double dx,
                           double dy, double* z, int64_t ldz, double* actual, int64_t lda,
                           int deriv_mode, double sx, double sy, int ghost_mask, double poison,
                           gmt_stream_t stream);

/* ===================================================================== fused
 * halo exchange + stencil kernels over peer memory.
 *
 * One (virtual) rank's view of an exchange.  A launch takes an array of these in DEVICE
 * memory and blockIdx.y selects the entry: a production process passes ONE entry whose
 * peer_* pointers are NVLink mappings of the neighbours' memory, while single-GPU
 * emulation passes every rank's entry in one cooperative launch (peers are then other
 * buffers on the same GPU).
 *
 * Protocol: each launch is epoch e = *epoch + 1.  A rank stores its boundary values into
 * the neighbour's inbox[e & 1] with ordinary stores, fences at system scope and publishes e
 * in the neighbour's flag word for that chunk (st.release.sys).  The consumer polls the
 * flag (ld.acquire.sys, bounded), copies the inbox into its own in-array ghost cells --
 * so after the kernel the array looks exactly as it does after the reference's
 * boundary_exchange -- and computes the boundary tiles last, interior tiles first.
 * A side with has == 0 is a physical boundary: its ghost cells are used as they are.    */
typedef struct {
  double* z;            /* n + 4 doubles, 32-byte aligned */
  double* out;          /* n doubles, 32-byte aligned */
  int64_t n;
  double scale;
  int32_t has[2];       /* 0 = low neighbour, 1 = high neighbour */
  double* my_inbox[2][2];   /* [side][parity] -> 2 doubles each, local memory */
  double* peer_inbox[2][2]; /* [side][parity] -> the neighbour's my_inbox[1-side][parity] */
  uint64_t* my_flag[2];     /* one word per side */
  uint64_t* peer_flag[2];   /* the neighbour's my_flag[1-side] */
  uint64_t* epoch;      /* device-resident epoch counter of this rank */
  uint32_t* done;       /* >= 2 zero-initialised words: CTA completion + tile counters (self-resetting) */
  uint32_t* status;     /* set non-zero when a bounded spin expired */
  void* err_scratch;    /* optional: gmt_reduce_scratch_bytes() of zeroed scratch */
  double* err_sumsq;    /* optional: receives sum((out - 3x^2)^2), x = x0 + i*dx */
  double x0, dx;
  int64_t spin_limit;   /* 0 = default */
} gmt_halo1d_rank;

/* replaces boundary_exchange + stencil1d_5 (mpi_stencil_gt.cc:83-122,206) */
int gmt_fused_stencil1d_f64(const gmt_halo1d_rank* ranks_dev, int n_ranks, int64_t n, int with_err,
                            gmt_stream_t stream);
/* The same step with the data path of gmt_stencil1d_tma_f64: windows streamed into a shared-memory
 * ring by cp.async.bulk (SASS UBLKCP) + mbarrier transaction counts; the producer thread also does the
 * halo export, the flag waits and the ghost landing (followed by fence.proxy.async before the bulk
 * copy that reads them).  No error norm.  Same descriptors / epoch: calls can be mixed. */
int gmt_fused_stencil1d_tma_f64(const gmt_halo1d_rank* ranks_dev, int n_ranks, int64_t n, gmt_stream_t stream);
/* The exchange alone -- boundary_exchange as a callable (mpi_stencil_gt.cc:83-122) and the
 * quantity the reference times (:200-204): store my two boundary values into each neighbour's
 * inbox, wait for theirs, put them into z's ghost cells.  Same descriptors, inboxes, flags and
 * epoch as gmt_fused_stencil1d_f64 (calls of both kinds can be mixed; each advances the epoch);
 * `out` and the error fields are not touched.  One 64-thread CTA per rank. */
int gmt_halo_exchange1d_f64(const gmt_halo1d_rank* ranks_dev, int n_ranks, gmt_stream_t stream);

/* 2-D: sides 0 = low x (dim 0), 1 = high x, 2 = low y (dim 1), 3 = high y.
 * gx, gy in {0, 2} are the ghost widths of z in each dim; z is (nx+2gx, ny+2gy) and
 * out(i,j) = sx * d/dx (if gx) + sy * d/dy (if gy).  gx=2,gy=0 is the reference's dim-0
 * slab test, gx=0,gy=2 its dim-1 test (boundary_exchange_x/_y + stencil2d_1d_5_d0/_d1,
 * mpi_stencil2d_gt.cc:136-373,530-532); gx=gy=2 is the 4-neighbour process-grid case of
 * BASELINE config 3.  Inbox layouts: x sides hold 2 doubles per column (inbox[2*j + r]),
 * y sides hold two contiguous ghost columns (inbox[r*nx + i]).                        */
typedef struct {
  double* z;
  double* out;
  int64_t nx, ny, ldz, ldo;
  int32_t gx, gy;
  double sx, sy;
  int32_t has[4];           /* 0: physical boundary (ghosts used as they are); 1: neighbour rank (export my
                               boundary to it, wait for its flags); 2: ghosts are FED BY THE HOST for this
                               step -- nothing is exported, the side's tiles wait for my_flag like for a
                               neighbour (the caller copies the values into my_inbox[side][e & 1] and then
                               sets the side's flag words to e, e.g. from a copy stream) */
  double* my_inbox[4][2];
  double* peer_inbox[4][2];
  uint64_t* my_flag[4];     /* gmt_halo2d_flag_count(side, nx, ny) words per side */
  uint64_t* peer_flag[4];
  uint64_t* epoch;
  uint32_t* done;
  uint32_t* status;
  void* err_scratch;
  double* err_sumsq;        /* optional: sum((out - (sx*3x^2*[gx] + sy*2y*[gy]))^2) */
  double x0, y0, dx, dy;
  int64_t spin_limit;
} gmt_halo2d_rank;

/* with_err: bit 0 = compute the error norm into err_sumsq; bit 1 (dim-0 slabs, gx == 2 && gy == 0,
 * only) = take the x-face ghost values straight from the inbox: `out` is the same, but z's own ghost
 * cells are NOT refreshed (no strided 16-byte writes into the array); call gmt_halo_exchange2d_f64
 * when the array itself must hold them. */
int gmt_fused_stencil2d_f64(const gmt_halo2d_rank* ranks_dev, int n_ranks, int64_t nx, int64_t ny,
                            int gx, int gy, int with_err, gmt_stream_t stream);
/* The exchange alone -- boundary_exchange_x / _y as a callable (mpi_stencil2d_gt.cc:136-373) and
 * the quantity the reference times (:512-522): every used side's two boundary rows / columns go
 * into the neighbour's inbox (strided 16-byte pieces gathered into contiguous peer stores for the
 * x faces, contiguous for the y faces), chunk flags follow, and what the neighbours (or the host,
 * has == 2) delivered is put into z's own ghost cells.  No sweep: afterwards z is what it is after
 * the reference's exchange, for any compute step of the caller's.  Same descriptors and epoch as
 * gmt_fused_stencil2d_f64; calls of both kinds can be mixed on one plan. */
int gmt_halo_exchange2d_f64(const gmt_halo2d_rank* ranks_dev, int n_ranks, int64_t nx, int64_t ny,
                            int gx, int gy, gmt_stream_t stream);
/* For sides with has == 2: after the host-to-device copy of the ghost values into my_inbox, publish
 * "arrived" for the step that is currently running / about to run: flags[i] <- *epoch + 1 (release).
 * Call it on the stream that carried the copy; it may run concurrently with that step's kernel. */
int gmt_publish_epoch_flags(uint64_t* flags, int64_t n, const uint64_t* epoch, gmt_stream_t stream);
/* Test hook: sleep a pseudo-random time of up to max_ns nanoseconds before every halo flag
 * store (0 = off, the default).  Shakes the flag/epoch/parity protocol in stress tests.   */
int gmt_debug_set_push_delay(unsigned int max_ns);
/* Test hook: `ctas` thread blocks of 1024 threads (two fill an SM) that hold their SM slots for
 * `ns` nanoseconds and do nothing else -- for tests that a kernel of this library does not depend on
 * having the whole GPU (no counterpart in the reference). */
int gmt_debug_occupy(int ctas, unsigned long long ns, gmt_stream_t stream);
/* Test hook: single-rank (non-cooperative) collective launches use up to `factor` times more CTAs
 * than can be resident at once (1 = off, the default).  A kernel that waited for a sibling CTA of its
 * own grid would then expire its bounded spin; the collectives here wait for other ranks only. */
int gmt_debug_set_grid_factor(int factor);
/* flag words needed on `side` for the given interior extents */
int64_t gmt_halo2d_flag_count(int side, int64_t nx, int64_t ny);

/* ================================================================ collectives
 * written as kernels over peer memory: no MPI, no NCCL.  Like the fused stencils they take a
 * DEVICE array of per-rank descriptors (blockIdx.y selects one): one entry per process in
 * production, all ranks in one cooperative launch when emulated on a single GPU.
 *
 * Signalling: every rank owns a flag block of GMT_COLL_FLAG_WORDS words; word [round*32 + src]
 * (rounds 0,1) holds the last epoch rank `src` signalled for that round (st.release.sys /
 * ld.acquire.sys, bounded spins).  Epochs only grow; reduction inputs travel through a staging
 * area that is double-buffered by epoch parity, so a rank that is one call ahead cannot disturb a
 * slower one.  Small payloads skip the flags altogether: they travel as 16-byte "LL" cells in the
 * staging area whose two 8-byte words each carry 32 payload bits and the 32-bit epoch, so data and
 * "arrived" come in one store (LL all-reduce, allgather of slices up to 512 KiB, and the per-CTA
 * partial sums of gmt_daxpy_allreduce_f64).
 *
 * Sizes: a staging area is gmt_coll_stage_bytes(world, stage_cap) bytes; stage_cap must be at least
 * gmt_coll_min_stage_cap() (= 2048: gmt_daxpy_allreduce_f64 stores one cell per CTA, up to 2048).
 * The raw entry points cannot see stage_cap from the host; a kernel that would need more staging
 * doubles or cells than the descriptor's stage_cap (all-reduce / axis-sum n, LL allgather slice,
 * daxpy partial cells) writes nothing, sets bit 0x1000 in the status word, still advances the
 * epoch, and (daxpy) returns NaN.  The engines and the Python layer check sizes on the host and return GMT_ETOOBIG. */
size_t gmt_coll_stage_bytes(int world, int64_t stage_cap);
int64_t gmt_coll_min_stage_cap(void);
int64_t gmt_coll_ll_gather_max(void);   /* largest slice (doubles) gmt_allgather_f64 takes in LL mode */
#define GMT_COLL_FLAG_WORDS 256
#define GMT_COLL_MAX_WORLD 32
typedef struct {
  int32_t rank, world;
  double** peer_buf;      /* device array [world]: each rank's symmetric USER buffer (allgather/gather) */
  double* mc_buf;         /* NVLS multicast address of that buffer, or NULL */
  double** peer_stage;    /* device array [world]: each rank's ZERO-INITIALISED staging area of
                             gmt_coll_stage_bytes(world, stage_cap) = (4 + 4*world)*stage_cap doubles: inputs
                             [parity][stage_cap], results [parity][stage_cap], then the "LL" cells
                             [parity][src rank][stage_cap] of 16 bytes */
  double* mc_stage;       /* multicast address of the staging area, or NULL */
  int64_t stage_cap;
  uint64_t** peer_flag;   /* device array [world]: each rank's flag block */
  uint64_t* epoch;        /* device-resident epoch counter of this rank */
  uint32_t* done;         /* >= 4 zero-initialised words (self-resetting) */
  uint32_t* status;       /* 1 + silent rank when a bounded spin expired; bit 0x1000: stage_cap too small for the call */
  void* scratch;          /* gmt_reduce_scratch_bytes() of zeroed local scratch (unused since the LL-cell daxpy) */
  const double* in;       /* local operand: allgather source / all-reduce input / axis-sum array / daxpy x */
  double* out;            /* local result: all-reduce output vector / global sum */
  double* y;              /* daxpy y */
  int64_t spin_limit;
} gmt_coll_rank;

/* in-place allgather: rank r's slice [r*n,(r+1)*n) is stored into the same offset of every
 * peer's buffer (MPI_Allgather(MPI_IN_PLACE), mpi_daxpy_nvtx.cc:285, mpigatherinplace.f90:39).
 * from_in != 0: the slice is read from `in` instead of the rank's own slot, and also stored
 * locally (MPI_Allgather with a separate send buffer, mpi_daxpy_nvtx.cc:288).
 * use_multicast (a mode): 0 = world-1 peer stores per 32 bytes; 1 = multimem.st.relaxed.sys through
 * mc_buf (one store per 16 bytes, the switch replicates); 3 = the same with weak multimem.st;
 * 4 = ordinary 256-bit stores on the multicast mapping; 2 = "LL": slices of up to
 * gmt_coll_ll_gather_max() doubles (and <= stage_cap) travel as epoch-carrying 16-byte cells through
 * the staging area and every rank copies them into its own buffer -- one one-way trip, no
 * signalling rounds, nobody writes anybody else's user buffer. */
int gmt_allgather_f64(const gmt_coll_rank* ranks_dev, int n_ranks, int64_t n, int from_in,
                      int use_multicast, gmt_stream_t stream);
/* gather to root in place (MPI_Gather with MPI_IN_PLACE at the root, BASELINE config 5).  Ends with an
 * all-to-all signalling round like every other collective here: no rank may get more than one call ahead
 * of another, or the parity double-buffering of the reductions' staging areas would not hold. */
int gmt_gather_f64(const gmt_coll_rank* ranks_dev, int n_ranks, int64_t n, int root, gmt_stream_t stream);
/* sum all-reduce of n doubles: out <- sum over ranks of in; out may alias in
 * (MPI_Allreduce(MPI_IN_PLACE), mpi_stencil2d_gt.cc:615,624).  n <= stage_cap.
 * algo: 0 one-shot (every rank reads every peer's vector), 1 two-shot (reduce-scatter then
 * allgather by peer stores), 2 one-shot through NVLS multimem.ld_reduce (needs mc_stage),
 * 3 "LL": every value is stored into every rank's staging area as a 16-byte cell that carries
 * the epoch in both of its 8-byte words (data and flag arrive together: no fence, no flag round
 * trip, no grid-wide rendezvous) -- the fastest up to a few hundred KiB.
 * Ranks add in rank order, so every rank gets the same bits (algos 0, 1 and 3).          */
int gmt_allreduce_f64(const gmt_coll_rank* ranks_dev, int n_ranks, int64_t n, int algo, gmt_stream_t stream);
/* fused daxpy + global sum: y <- a*in + y on each rank, every rank's *out <- sum over all ranks
 * of sum(y_new)  (cublasDaxpy + MPI_Allgather + host sums, mpi_daxpy_nvtx.cc:245-310, in one kernel).
 * One-shot grid in address order; every CTA leaves its partial as an LL cell in its own rank's
 * staging area, the last CTA adds them in a fixed order, and the rank totals cross NVLink as one LL
 * cell per peer and are added in rank order (same bits on every rank, reproducible).  Needs
 * stage_cap >= gmt_coll_min_stage_cap(); n must be the same on every rank. */
int gmt_daxpy_allreduce_f64(const gmt_coll_rank* ranks_dev, int n_ranks, int64_t n, double a,
                            gmt_stream_t stream);
/* axis sum of the column-major (nx, ny, ld) array `in` whose epilogue all-reduces the first n_out
 * entries across ranks into `out` (gt::sum_axis_to + MPI_Allreduce, mpi_stencil2d_gt.cc:611-625).
 * n_out <= stage_cap.  LL cells while n_out x world <= 2^18, one-shot beyond (chosen in the kernel,
 * which knows the world size).  Safe without co-residency of its grid: whoever waits for a cell
 * waits for another rank only. */
int gmt_axis_sum_allreduce_f64(const gmt_coll_rank* ranks_dev, int n_ranks, int64_t nx, int64_t ny,
                               int64_t ld, int axis, int64_t n_out, gmt_stream_t stream);
/* signal barrier across ranks (MPI_Barrier, mpi_daxpy_nvtx.cc:277) */
int gmt_barrier(const gmt_coll_rank* ranks_dev, int n_ranks, gmt_stream_t stream);

/* ============================================================ native exchange engine
 * (engine.cu) A whole px x py process grid driven from ONE C/C++ process that owns all GPUs
 * of the node -- the C++ counterpart of the reference's per-program host orchestration
 * (neighbours, ghosted arrays, exchange + stencil per iteration, timing:
 * mpi_stencil2d_gt.cc:385-572) with no Python and no MPI.  Rank r sits at (r % px, r / px),
 * on devices[r] or, with devices == NULL, on the device the reference's block rule picks
 * (set_rank_device, mpi_daxpy.cc:36-62: when there are more ranks than GPUs, consecutive
 * ranks share one).  Neighbouring ranks on different GPUs are wired through CUDA peer
 * access; ranks sharing a GPU go into one cooperative launch.  Everything is allocated in
 * create(); step() only launches (one fused halo+stencil kernel per GPU, asynchronous).
 * Per-rank interior is nx x ny (weak scaling), domain [0,8) per decomposed direction,
 * z = x^3 + y^2, out = d/dx (gx == 2) + d/dy (gy == 2).  Slab decompositions (only one of gx, gy)
 * use ONE spacing, 8 / n_global of the differentiated axis, for both coordinates, like the
 * reference's test_deriv (mpi_stencil2d_gt.cc:427-448).  Not thread-safe per engine.    */
typedef struct gmt_engine2d gmt_engine2d;
int gmt_engine_default_device(int n_ranks, int rank, int n_devices, int* device);
int gmt_engine2d_create(int px, int py, const int* devices, int64_t nx, int64_t ny, int gx, int gy,
                        int with_err, int64_t spin_limit, gmt_engine2d** out);
int gmt_engine2d_destroy(gmt_engine2d* e);
int gmt_engine2d_world(const gmt_engine2d* e);
int gmt_engine2d_device_groups(const gmt_engine2d* e);
int gmt_engine2d_rank_device(const gmt_engine2d* e, int rank, int* device);
/* which: 0 = ghosted input z, 1 = result out.  Column-major: `cols` columns of leading dimension ld. */
int gmt_engine2d_array(const gmt_engine2d* e, int rank, int which, double** ptr, int64_t* cols, int64_t* ld);
/* analytic field; ghosts analytic on physical sides, `poison` on sides a neighbour must fill */
int gmt_engine2d_init_analytic(gmt_engine2d* e, double poison);
int gmt_engine2d_step(gmt_engine2d* e);          /* asynchronous */
/* the halo exchange alone (gmt_halo_exchange2d_f64 on every GPU), asynchronous; can be mixed with step() */
int gmt_engine2d_exchange(gmt_engine2d* e);
int gmt_engine2d_synchronize(gmt_engine2d* e);
/* GMT_ETIMEOUT (and the rank) when some rank's bounded flag wait expired */
int gmt_engine2d_status(gmt_engine2d* e, int* first_bad_rank);
/* warm-up, then `steps` steps bracketed by CUDA events on every GPU's stream; max over GPUs */
int gmt_engine2d_time_steps(gmt_engine2d* e, int steps, int warmup, double* ms_per_step);
/* the same for exchange-only calls: the reference's timed quantity (mpi_stencil2d_gt.cc:512-522) */
int gmt_engine2d_time_exchanges(gmt_engine2d* e, int steps, int warmup, double* ms_per_exchange);
/* sqrt of the in-kernel sum((out - analytic)^2) of the last step (needs with_err) */
int gmt_engine2d_err_norm(gmt_engine2d* e, int rank, double* err_norm);

/* Collective engine (engine_coll.cu): the same single-process arrangement for the collectives --
 * the host side of the reference's daxpy benchmark and gather test (mpi_daxpy_nvtx.cc:85-343,
 * mpigatherinplace.f90:1-58).  Every rank owns a symmetric user buffer of buf_doubles doubles
 * (gathered arrays live there) and a staging area of gmt_coll_stage_bytes(world, stage_cap) bytes
 * (stage_cap is raised to gmt_coll_min_stage_cap() if smaller); each call below is one
 * asynchronous launch per GPU on the engine's per-GPU stream (gmt_coll_engine_stream: order your
 * own work on a rank's operands on that stream).  No NVLS multicast variants here.            */
typedef struct gmt_coll_engine gmt_coll_engine;
int gmt_coll_engine_create(int world, const int* devices, int64_t buf_doubles, int64_t stage_cap,
                           int64_t spin_limit, gmt_coll_engine** out);
int gmt_coll_engine_destroy(gmt_coll_engine* e);
int gmt_coll_engine_world(const gmt_coll_engine* e);
int gmt_coll_engine_device_groups(const gmt_coll_engine* e);
int gmt_coll_engine_rank_device(const gmt_coll_engine* e, int rank, int* device);
int gmt_coll_engine_buffer(const gmt_coll_engine* e, int rank, double** ptr, int64_t* doubles);
int gmt_coll_engine_stream(const gmt_coll_engine* e, int rank, gmt_stream_t* stream);
/* local operands of a rank for the next calls (see gmt_coll_rank.in/out/y); stream-ordered */
int gmt_coll_engine_bind(gmt_coll_engine* e, int rank, const double* in, double* out, double* y);
int gmt_coll_engine_allgather(gmt_coll_engine* e, int64_t n, int from_in);
int gmt_coll_engine_gather(gmt_coll_engine* e, int64_t n, int root);
int gmt_coll_engine_allreduce(gmt_coll_engine* e, int64_t n, int algo);
int gmt_coll_engine_daxpy_allreduce(gmt_coll_engine* e, int64_t n, double a);
int gmt_coll_engine_axis_sum_allreduce(gmt_coll_engine* e, int64_t nx, int64_t ny, int64_t ld, int axis,
                                       int64_t n_out);
int gmt_coll_engine_barrier(gmt_coll_engine* e);
int gmt_coll_engine_synchronize(gmt_coll_engine* e);
int gmt_coll_engine_status(gmt_coll_engine* e, int* first_bad_rank);
/* device timing: mark(0) / mark(1) record an event on every GPU's stream; elapsed_ms waits for
 * the second marks and returns the largest per-GPU interval */
int gmt_coll_engine_mark(gmt_coll_engine* e, int which);
int gmt_coll_engine_elapsed_ms(gmt_coll_engine* e, double* ms);

#ifdef __cplusplus
}
#endif
#endif /* GMT_B200_H */
