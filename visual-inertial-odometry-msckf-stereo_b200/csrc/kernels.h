// Host-side launchers of the CUDA kernels (one translation unit per kernel family).
#pragma once
#include <cuda_runtime.h>
#include <atomic>
#include <mutex>
#include <vector>
#include "msckf_dev.cuh"

namespace msckf {

// One-time work per device (cudaFuncSetAttribute is per device; a process may hold contexts on several GPUs and
// drive them from several host threads).
struct DevOnce {
  std::atomic<unsigned long long> done{0};
  std::mutex mu;
  template <class Fn>
  void run(Fn&& fn) {
    int dev = 0;
    cudaGetDevice(&dev);
    const unsigned long long bit = 1ull << (dev & 63);
    if (done.load(std::memory_order_acquire) & bit) return;
    std::lock_guard<std::mutex> g(mu);
    if (done.load(std::memory_order_relaxed) & bit) return;
    fn();
    done.fetch_or(bit, std::memory_order_release);
  }
};

// Kernel ids for launch counting and optional per-kernel event timing (msckf_kernel_times).
enum KernelId {
  KID_PROPAGATE = 0, KID_AUGMENT, KID_CAM_PREPARE, KID_OBSERVE, KID_CLASSIFY_LOST, KID_PRUNE_SELECT, KID_TRIANGULATE, KID_PLAN_ITEMS,
  KID_JACOBIAN_GATE, KID_JACOBIAN_ROWS, KID_JACOBIAN_BIG, KID_PLAN_STACK, KID_QR, KID_QR_WIDE, KID_QR_TALL, KID_QR_D16, KID_QR_D32, KID_QR_D48, KID_QR_D64, KID_QR_D80, KID_UPD_DMMA16, KID_UPD_DMMA32, KID_UPD_DMMA48, KID_UPD_DMMA64, KID_UPD_DMMA80, KID_UPD_T, KID_UPD_S, KID_UPD_CHOL, KID_UPD_TRSM, KID_UPD_INJECT,
  KID_UPD_DOWNDATE, KID_REMOVE_CAMS, KID_FINISH_FRAME, KID_RESET, KID_MISC, KID_COUNT
};
extern const char* const kKernelNames[KID_COUNT];

// Host-side launch bookkeeping of one context.  Counting is always on; event timing only when
// `timing` is set (two events per launch on the launching stream, accumulated at flush()).
struct KTimer {
  bool timing = false;
  int bad_launch = -1;           // kernel id of the first launch the runtime rejected (reported with the error text)
  long long launches[KID_COUNT] = {};
  double ms[KID_COUNT] = {};
  long long timed[KID_COUNT] = {};
  std::vector<cudaEvent_t> pool;
  std::vector<int> pending;      // kernel id of every recorded (start, stop) pair
  void begin(int id, cudaStream_t st);
  void end(cudaStream_t st);
  void flush(cudaStream_t st);
  void release();
};
#define KLAUNCH(d, id, st, ...)                         \
  do {                                                  \
    KTimer* kt_ = static_cast<KTimer*>((d).host_timer); \
    if (kt_) kt_->begin((id), (st));                    \
    __VA_ARGS__;                                        \
    if (kt_ && kt_->bad_launch < 0 && cudaPeekAtLastError() != cudaSuccess) kt_->bad_launch = (id); \
    if (kt_) kt_->end((st));                            \
  } while (0)

// Independent launches of one step (the size classes of the Jacobian, QR and update kernels work on disjoint
// landmarks / filters) are spread over auxiliary streams of the context and joined back into its main stream.
constexpr int N_AUX = 4;
struct Lanes {
  cudaStream_t aux[N_AUX] = {};
  cudaEvent_t fork = nullptr, join[N_AUX] = {}, built = nullptr;
  bool enabled = false;
};
struct Fork {
  cudaStream_t main;
  Lanes* L;
  bool on;
  unsigned used = 0;
  Fork(const Dev& d, cudaStream_t st) : main(st), L(static_cast<Lanes*>(d.host_lanes)) {
    const KTimer* kt = static_cast<const KTimer*>(d.host_timer);
    on = L && L->enabled && !(kt && kt->timing);          // the per-kernel event pass runs everything in order
    if (on) cudaEventRecord(L->fork, main);
  }
  cudaStream_t lane(int i) {                               // lane 0 is the main stream
    if (!on || i <= 0) return main;
    i = (i - 1) % N_AUX;
    if (!((used >> i) & 1u)) {
      cudaStreamWaitEvent(L->aux[i], L->fork, 0);
      used |= 1u << i;
    }
    return L->aux[i];
  }
  ~Fork() {
    if (!on) return;
    for (int i = 0; i < N_AUX; ++i)
      if ((used >> i) & 1u) {
        cudaEventRecord(L->join[i], L->aux[i]);
        cudaStreamWaitEvent(main, L->join[i], 0);
      }
  }
};

// k_propagate.cu
void launch_propagate(const Dev& d, const int* imu_off, const double* ts, const double* gyro, const double* acc, cudaStream_t st,
                      const FrameIn* fin = nullptr);
void launch_cam_slot(const Dev& d, cudaStream_t st);
void launch_augment(const Dev& d, const long long* ids, cudaStream_t st, const FrameIn* fin = nullptr, bool slot_done = false);
void launch_finish_frame(const Dev& d, int do_reset_check, cudaStream_t st);
void launch_reset(const Dev& d, int only_filter, int reset_imu, cudaStream_t st);
void launch_remove_cams(const Dev& d, cudaStream_t st);

// k_store.cu  (landmark store = map_server bookkeeping)
void launch_observe(const Dev& d, const int* feat_off, const long long* ids, const double* uv, cudaStream_t st,
                    const FrameIn* fin = nullptr, int pending_augment = 0);
void launch_classify_lost(const Dev& d, cudaStream_t st);
void launch_plan_items(const Dev& d, int stage, cudaStream_t st);
void launch_plan_stack(const Dev& d, int stage, cudaStream_t st);
void launch_prune_select(const Dev& d, cudaStream_t st);
void launch_find_redundant(const Dev& d, int filter, long long* out_ids, cudaStream_t st);

// k_triangulate.cu
void launch_triangulate_store(const Dev& d, cudaStream_t st);
void launch_triangulate_raw(const Dev& d, int n, const int* obs_off, const double* uv, const double* cam_q, const double* cam_p,
                            double* pos, unsigned char* valid, int* solves, int max_obs, int mode, cudaStream_t st);

// k_jacobian.cu
void launch_jacobian_gate(const Dev& d, int class_mask, int with_big, int write_all, cudaStream_t st);
int sm_count();
void launch_jacobian_rows(const Dev& d, int cls, int write_all, cudaStream_t st);   // k_jacobian_rows.cu (lane-per-row kernel, short tracks)
void launch_cam_prepare(const Dev& d, cudaStream_t st);
void launch_gate_lu(const Dev& d, cudaStream_t st);
void launch_measurement_jacobian(const Dev& d, int filter, int cam_logical, int slot, double* out /*[4*6+4*3+4]*/, cudaStream_t st);
void launch_export_item_dense(const Dev& d, int filter, double* H, double* r, int n_state, cudaStream_t st);
void launch_gate_dense(const Dev& d, int filter, const double* H, const double* r, int m, int n_state, int dof,
                       double* scratch, double* out_gamma, int* out_pass, cudaStream_t st);

// k_update.cu
void launch_qr_build(const Dev& d, int stage, cudaStream_t st);   // wide / tall / uncompressed stacks
void launch_qr_wide(const Dev& d, cudaStream_t st);               // k_qr_wide.cu: TSQR over the grid
int qr_wide_rows_per_chunk(int Nmax);
void launch_qr_dmma(const Dev& d, Fork& fk);               // k_qr_dmma.cu
void launch_qr_dmma_class(const Dev& d, int cls, cudaStream_t st);
void launch_upd_dmma_class(const Dev& d, int cls, cudaStream_t st, int stage = 0);   // k_update_dmma.cu (stage 1: compacted output)
void launch_compress_update(const Dev& d, int stage, cudaStream_t st); // QR + update of every class, each class on its own stream
void launch_qr_tall(const Dev& d, cudaStream_t st);                    // k_qr_tall.cu
void launch_qr_dense(const Dev& d, int filter, int mp_dense, cudaStream_t st);
void launch_update(const Dev& d, int stage, cudaStream_t st);
void launch_update_tiled(const Dev& d, cudaStream_t st);
void launch_update_dmma(const Dev& d, Fork& fk);             // k_update_dmma.cu
size_t upd_dmma_smem_bytes(int nmax, int KC);
void launch_load_dense_H(const Dev& d, int filter, const double* H, const double* r, int m, int n_state, cudaStream_t st);

}  // namespace msckf
