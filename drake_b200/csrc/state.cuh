// state.cuh -- host-side owner of the device state and the launch logic of one MPM scene.
// Replaces GpuMpmState<T> (cuda_mpm_model.cuh/.cu) and GpuMpmSolver<T> (cuda_mpm_solver.cu) of the
// reference; exposed through the C ABI in gmpm_capi.cu.  No CPU compute path exists here.
#pragma once
#include <algorithm>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "../../include/gmpm_c.h"
#include "common.cuh"
#include "halo.cuh"
#include "kernels.cuh"
#include "p2g_col.cuh"
#include "p2g_pair.cuh"
#include "volumetric.cuh"
#include "sort.cuh"

namespace gmpm {

// Event-based per-kernel timing: a mark is recorded on the stream before every launch; the time
// between two consecutive marks is attributed to the first one's kernel.
struct Profiler {
  bool on = false;
  std::vector<cudaEvent_t> pool;
  std::vector<std::pair<const char*, cudaEvent_t>> marks;
  std::map<std::string, std::pair<double, uint64_t>> acc;
  size_t next = 0;
  cudaEvent_t get() {
    if (next == pool.size()) {
      cudaEvent_t e;
      GMPM_CUDA(cudaEventCreate(&e));
      pool.push_back(e);
    }
    return pool[next++];
  }
  void mark(const char* name, cudaStream_t st) {
    cudaEvent_t e = get();
    GMPM_CUDA(cudaEventRecord(e, st));
    marks.emplace_back(name, e);
    if (marks.size() >= 4096) flush(st);
  }
  void flush(cudaStream_t st) {
    if (marks.empty()) return;
    cudaEvent_t e = get();
    GMPM_CUDA(cudaEventRecord(e, st));
    GMPM_CUDA(cudaEventSynchronize(e));
    for (size_t i = 0; i < marks.size(); ++i) {
      cudaEvent_t b = marks[i].second, n = (i + 1 < marks.size()) ? marks[i + 1].second : e;
      float ms = 0.f;
      GMPM_CUDA(cudaEventElapsedTime(&ms, b, n));
      auto& a = acc[marks[i].first];
      a.first += ms;
      a.second += 1;
    }
    marks.clear();
    next = 0;
  }
  void reset() {
    marks.clear();
    acc.clear();
    next = 0;
  }
  ~Profiler() {
    for (auto e : pool) cudaEventDestroy(e);
  }
};

enum class GridStage { kClean, kTiles, kMomentum, kVelocity };
template <typename T> struct ContactScalars;  // contact.cuh

// SyncParticleStateToCpu with declared query bounds (gmpm_set_query_bounds): positions in the reference's layout and
// visible order go to pos_stage as always (device-side consumers), and the particles inside at least one of the
// axis-aligned boxes are appended -- visible index + position -- to the candidate list that crosses PCIe.
constexpr int kMaxQueryBounds = 64;
template <typename T>
struct QueryBounds {
  T lo[kMaxQueryBounds][3], hi[kMaxQueryBounds][3];
  int n;
};
template <typename T>
__global__ void cull_positions_kernel(uint32_t N, const T* __restrict__ src, const int* __restrict__ pid_of_vis,
                                      const int* __restrict__ slot_of_pid, T* __restrict__ pos_stage,
                                      const QueryBounds<T>* __restrict__ qb, uint32_t* __restrict__ count,
                                      uint32_t* __restrict__ cand_idx, T* __restrict__ cand_pos) {
  const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
  bool hit = false;
  T x[3] = {T(0), T(0), T(0)};
  if (v < N) {
    const int s = slot_of_pid[pid_of_vis[v]];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      x[c] = src[(size_t)c * N + s];
      pos_stage[(size_t)v * 3 + c] = x[c];
    }
    const int n = qb->n;
    for (int b = 0; b < n && !hit; ++b)
      hit = x[0] >= qb->lo[b][0] && x[0] <= qb->hi[b][0] && x[1] >= qb->lo[b][1] && x[1] <= qb->hi[b][1] &&
            x[2] >= qb->lo[b][2] && x[2] <= qb->hi[b][2];
  }
  const uint32_t m = __ballot_sync(0xFFFFFFFFu, hit);
  if (!m) return;
  const int lane = threadIdx.x & 31;
  uint32_t base = 0;
  if (lane == __ffs(m) - 1) base = atomicAdd(count, (uint32_t)__popc(m));
  base = __shfl_sync(0xFFFFFFFFu, base, __ffs(m) - 1);
  if (hit) {
    const uint32_t o = base + __popc(m & ((1u << lane) - 1u));
    cand_idx[o] = v;
#pragma unroll
    for (int c = 0; c < 3; ++c) cand_pos[(size_t)o * 3 + c] = x[c];
  }
}

template <typename T>
struct State {
  gmpm_params prm{};
  Consts<T> k{};
  int device = 0;
  cudaStream_t stream = nullptr;
  bool finalized = false;

  // host staging (AddQRCloth, cuda_mpm_model.cu:16-33)
  std::vector<T> h_pos, h_vel;
  std::vector<int> h_tri;
  size_t n_verts = 0, n_faces = 0, N = 0;
  // volumetric particles (gmpm_add_particles; pids [n_faces + n_verts, N)), SURVEY 8f-4
  size_t n_vol = 0;
  std::vector<T> h_vpos, h_vvel, h_vvol;
  std::vector<uint8_t> h_vmat;
  std::vector<gmpm_material> materials;
  T* Fv = nullptr;            // [n_vol][9] row-major, by volumetric particle id
  uint8_t* mat_id = nullptr;  // [n_vol]
  MaterialTable<T> mat_table{};
  uint32_t n_cloth() const { return (uint32_t)(n_faces + n_verts); }

  // particles (physical slots), double buffered
  T *pos[2]{}, *vel[2]{}, *C[2]{}, *vol[2]{};
  int* pid[2]{};
  int cur = 0;
  T *forces = nullptr, *taus = nullptr;
  int* slot_of_pid = nullptr;
  // faces
  // face records, double buffered for the spatial re-ordering (fid[j] = face id of record j)
  int *tri = nullptr, *fid = nullptr, *tri_alt = nullptr, *fid_alt = nullptr;
  T *Fd = nullptr, *Dm_inv = nullptr, *Fd_alt = nullptr, *Dm_inv_alt = nullptr;
  uint32_t* face_keys = nullptr;
  uint64_t fem_calls = 0;
  // visible order (reference semantics: current_pids / index_mappings / sort_keys / sort_ids)
  int* pid_of_vis[2]{};
  int vis_cur = 0;
  int* vis_of_pid = nullptr;
  uint32_t *vis_keys[2]{}, *vis_ids[2]{};
  int vis_res = 0;
  bool vis_keys_valid = false;
  // binning
  uint32_t* key = nullptr;
  uint32_t *skeys[2]{}, *svals[2]{};
  int sres = 0;
  uint32_t* sort_scratch = nullptr;
  uint32_t *pb_start = nullptr, *pb_end = nullptr, *pb_epoch = nullptr, *pb_list = nullptr;
  uint2* pb_tile = nullptr;
  WorkRec* pb_work = nullptr;
  size_t work_cap = 0;
  bool keys_valid = false, bin_valid = false, fem_valid = false;
  uint32_t bin_epoch = 0, p2g_epoch = 0;
  // grid
  T *tiles = nullptr, *g_mass = nullptr, *g_vel = nullptr, *g_vstar = nullptr;
  uint32_t* touched_epoch = nullptr;
  uint32_t* touched_list[2]{};
  int tl_cur = 0;
  GridStage stage = GridStage::kClean;
  DevCounters* ctr = nullptr;
  // contact (contact.cuh)
  size_t n_contacts = 0, contact_cap = 0;
  uint32_t *c_pid = nullptr, *c_rigid_id = nullptr, *c_keys = nullptr;
  T *c_pos = nullptr, *c_vel = nullptr, *c_vel0 = nullptr, *c_dist = nullptr, *c_normal = nullptr,
    *c_rigid_v = nullptr, *c_rigid_p = nullptr, *c_frame = nullptr;
  // contact blocks (grid blocks a contact stencil reaches) and the compact Hess/Grad/Dir store
  uint32_t *cb_epoch = nullptr, *cb_slot = nullptr, *cb_list = nullptr;
  size_t cb_cap = 0;
  uint32_t contact_epoch = 0;
  bool cb_valid = false;
  T *g_Hess = nullptr, *g_Grad = nullptr, *g_Dir = nullptr;
  size_t n_bodies = 0, body_cap = 0;
  T *F_tau = nullptr, *F_f = nullptr;
  ContactScalars<T>* csc = nullptr;  // device scalars of the contact solve
  void* h_csc = nullptr;             // pinned mirror (stats read-back)
  void* h_stage = nullptr;           // pinned staging of CopyContactPairs
  size_t h_stage_cap = 0;
  int newton_ctas = 0;
  // io
  void* io_buf = nullptr;
  size_t io_cap = 0;
  void* h_pinned = nullptr;
  size_t h_pinned_cap = 0;
  // asynchronous positions_host() mirror: device staging (visible order, reference layout), pinned host
  // mirror, a copy stream and two events (see positions_async)
  T* pos_stage = nullptr;
  T* h_pos_mirror = nullptr;
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t ev_gather = nullptr, ev_copy = nullptr;
  bool copy_pending = false;
  // culled snapshot (gmpm_set_query_bounds)
  bool cull_on = false, cull_pending = false;
  QueryBounds<T>* d_qb = nullptr;
  uint32_t* cand_count = nullptr;   // device
  uint32_t* cand_idx = nullptr;
  T* cand_pos = nullptr;
  uint32_t* h_cand_count = nullptr; // pinned
  uint32_t* h_cand_idx = nullptr;
  T* h_cand_pos = nullptr;
  std::vector<uint32_t> h_prev_idx; // mirror entries that currently hold a real position
  cudaEvent_t ev_count = nullptr;
  uint32_t cand_last = 0, cand_fetched = 0;  // candidates of the last snapshot / fetched speculatively this time
  uint64_t snapshot_d2h_bytes = 0;  // bytes the last snapshot moved to the host (instrumentation)
  bool driver_geometry_frozen = false;  // run_driver_substeps: keep the rigid poses for all substeps (reference driver)
  bool snapshot_valid = false;  // pos_stage holds the positions of the current substep (cleared by g2p)
  // device-side contact generation (contact_gen.cuh)
  uint32_t* gen_counts = nullptr;
  unsigned char* gen_geoms = nullptr;
  uint32_t gen_cap = 0;

  // shared-block exchange (halo.cuh)
  uint8_t *halo_flags_d = nullptr, *halo_own = nullptr;
  uint8_t* halo_flags_buf[2]{};   // the plans alternate between two flag buffers; the idle one is zeroed on the side stream
  bool halo_flags_clean[2]{};
  int halo_flags_idx = 0;
  uint8_t* halo_mark = nullptr;   // 1 for the blocks of the exchange in flight (part-wise UpdateGrid / G2P)
  uint32_t* halo_nbr = nullptr;   // per touched-list entry: outer-layer occupancy towards the 26 neighbours
  uint32_t halo_nbr_epoch = 0;    // P2G epoch halo_nbr was computed for
  uint32_t halo_nbr_zeroed_epoch = 0;
  uint32_t *halo_list[2]{}, *halo_cta_counts = nullptr, *halo_dev_words = nullptr;  // [0..1] counts, [2] error
  uint32_t* h_halo_words = nullptr;  // pinned mirror of halo_dev_words (per plan slot: count, error)
  T* halo_packed = nullptr;
  size_t halo_cap = 0;            // shared blocks the lists / packed buffer hold
  uint32_t halo_n = 0;            // shared blocks of the current exchange
  int halo_cur = 0;               // list used by the current exchange; the next plan goes to halo_cur ^ 1
  bool halo_planned = false, halo_flags_out = false, halo_own_valid = false, halo_packed_out = false;
  cudaStream_t halo_stream = nullptr;
  cudaEvent_t ev_halo_flags = nullptr, ev_halo_plan = nullptr;
  // peer mode (halo.cuh): exchange through receive windows in peer memory instead of an all-reduce
  bool halo_peer = false;
  uint32_t halo_rank = 0, halo_world = 1, halo_win_cap = 0, halo_xchg = 0;
  unsigned char* halo_window = nullptr;          // this rank's window (plain cudaMalloc: exportable through CUDA IPC)
  size_t halo_window_bytes = 0;
  HaloWindows halo_windows{};                    // every rank's window as mapped into this process
  bool halo_window_ipc[kHaloMaxPeers]{};         // opened with cudaIpcOpenMemHandle (to be closed)
  uint8_t* halo_sharers[2]{};                    // sharer set of each listed block
  uint32_t* halo_index[2]{};                     // grid block -> list position, per plan slot (fused exchange)
  bool halo_fused = std::getenv("GMPM_HALO_UNFUSED") == nullptr;  // send in the grid assembly, receive in the update
  bool halo_recv_pending = false;                // halo_peer_recv was called: the next UpdateGrid receives
  bool halo_sent = false;
  bool fem_reverse = std::getenv("GMPM_FEM_FORWARD") == nullptr;  // fem_kernel: walk the face records from the end

  std::vector<void*> allocs;
  size_t device_bytes = 0;
  uint64_t launches = 0;
  Profiler prof;
  int sm_count = 148;
  static constexpr int kP2GDefault = 0;  // 0 tile (default: fastest at fp32, within 3 % at fp64), 1 column, 2 column + bulk tile store (profiles/r02_p2g_ab.txt)
  int p2g_variant = kP2GDefault;
  int p2g_threads = kP2GThreads;
  size_t p2g_smem = 0;
  int p2g_ctas_per_sm = 1, g2p_ctas_per_sm = 1, fem_ctas_per_sm = 1, bin_ctas = 0, sort_ctas = 0;
  bool g2p_queue_ready = false;
  bool forces_zero = true;      // forces[] is all zeros (fresh allocation, or g2p just wrote the zeros)
  bool prologue_done = false;   // the binning kernel already rolled the touched list for the coming P2G

  // ---------------------------------------------------------------------------------------------
  template <typename U>
  U* dalloc(size_t n, bool zero = true) {
    void* p = nullptr;
    const size_t bytes = std::max<size_t>(n, 1) * sizeof(U);
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e != cudaSuccess) {
      char buf[160];
      snprintf(buf, sizeof(buf), "cudaMalloc(%zu bytes) failed: %s", bytes, cudaGetErrorString(e));
      throw CudaError(e, buf);
    }
    allocs.push_back(p);
    device_bytes += bytes;
    if (zero) GMPM_CUDA(cudaMemsetAsync(p, 0, bytes, stream));
    return static_cast<U*>(p);
  }
  void launch(const char* name) {
    ++launches;
    if (prof.on) prof.mark(name, stream);
  }
  void note(const char* name) {  // non-kernel stream work (memset / memcpy) for the profiler only
    if (prof.on) prof.mark(name, stream);
  }
  static uint32_t blocks_for(size_t n, int threads = kThreads) { return (uint32_t)((n + threads - 1) / threads); }

  void init(const gmpm_params& p) {
    prm = p;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) throw CudaError(cudaErrorNoDevice, "no CUDA device available (libgmpm_b200 has no CPU fallback)");
    if (p.device >= 0) GMPM_CUDA(cudaSetDevice(p.device));
    GMPM_CUDA(cudaGetDevice(&device));
    GMPM_CUDA(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, device));
    GMPM_CUDA(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    if (p.domain_bits < 3 || p.domain_bits > 10) throw std::invalid_argument("domain_bits must be in [3,10]");
    if (p.gravity_axis < 0 || p.gravity_axis > 2) throw std::invalid_argument("gravity_axis must be 0,1,2");
    k.domain_bits = p.domain_bits;
    k.domain_size = 1 << p.domain_bits;
    k.n_cells = 1u << (3 * p.domain_bits);
    k.n_blocks = 1u << (3 * (p.domain_bits - kBlockBits));
    k.boundary_cells = p.boundary_cells;
    k.gravity_axis = p.gravity_axis;
    if (p.dx < 0) throw std::invalid_argument("dx must be >= 0");
    if (p.dx > 0) {                           // cell size given: domain = [0, dx * 2^bits]^3
      k.dx = T(p.dx);
      k.dx_inv = T(1) / k.dx;
    } else {
      k.dx_inv = T(1) * T(1 << p.domain_bits);  // settings.h:50,57
      k.dx = T(1) / k.dx_inv;
    }
    k.d_inv = T(4) * k.dx_inv * k.dx_inv;     // settings.h:59
    const T E = T(p.youngs), nu = T(p.poisson);
    k.mu = E / (T(2) * (T(1) + nu));          // settings.h:114-115
    k.lambda = E * nu / ((T(1) + nu) * (T(1) - T(2) * nu));
    k.density = T(p.density);
    k.gamma = T(p.gamma);
    k.k_normal = T(p.k_normal);
    k.v_blend = T(p.v_blend);
    k.c_f = T(p.c_f);
    k.sdf_friction = T(p.sdf_friction);
    k.gravity = T(p.gravity);
    k.epsv = T(p.epsv);
  }

  void destroy() {
    if (bin_stream) {
      cudaStreamSynchronize(bin_stream);
      cudaStreamDestroy(bin_stream);
      cudaEventDestroy(ev_prekeys);
      cudaEventDestroy(ev_bin_done);
      bin_stream = nullptr;
      bin_async_pending = false;
    }
    if (stream) cudaStreamSynchronize(stream);
    for (void* p : allocs) cudaFree(p);
    allocs.clear();
    if (io_buf) cudaFree(io_buf);
    if (h_pinned) cudaFreeHost(h_pinned);
    if (h_csc) cudaFreeHost(h_csc);
    if (h_stage) cudaFreeHost(h_stage);
    if (halo_stream) {
      cudaStreamSynchronize(halo_stream);
      cudaStreamDestroy(halo_stream);
      halo_stream = nullptr;
    }
    for (int q = 0; q < kHaloMaxPeers; ++q)
      if (halo_window_ipc[q] && halo_windows.base[q]) {
        cudaIpcCloseMemHandle(halo_windows.base[q]);
        halo_window_ipc[q] = false;
      }
    if (halo_window) cudaFree(halo_window);
    halo_window = nullptr;
    halo_windows = HaloWindows{};
    if (h_halo_words) cudaFreeHost(h_halo_words);
    h_halo_words = nullptr;
    if (ev_halo_flags) cudaEventDestroy(ev_halo_flags);
    if (ev_halo_plan) cudaEventDestroy(ev_halo_plan);
    ev_halo_flags = ev_halo_plan = nullptr;
    if (copy_stream) {
      cudaStreamSynchronize(copy_stream);
      cudaStreamDestroy(copy_stream);
      copy_stream = nullptr;
    }
    if (h_pos_mirror) cudaFreeHost(h_pos_mirror);
    h_pos_mirror = nullptr;
    if (h_cand_count) cudaFreeHost(h_cand_count);
    if (h_cand_idx) cudaFreeHost(h_cand_idx);
    if (h_cand_pos) cudaFreeHost(h_cand_pos);
    h_cand_count = h_cand_idx = nullptr;
    h_cand_pos = nullptr;
    d_qb = nullptr;  // device buffers are released with the allocation list
    cull_on = cull_pending = false;
    if (ev_count) cudaEventDestroy(ev_count);
    ev_count = nullptr;
    if (ev_gather) cudaEventDestroy(ev_gather);
    if (ev_copy) cudaEventDestroy(ev_copy);
    ev_gather = ev_copy = nullptr;
    copy_pending = false;
    io_buf = nullptr;
    h_pinned = nullptr;
    h_csc = nullptr;
    h_stage = nullptr;
    for (auto& t : timers)
      if (t) {
        cudaEventDestroy(t);
        t = nullptr;
      }
    if (stream) cudaStreamDestroy(stream);
    stream = nullptr;
  }

  void bind_device() const { GMPM_CUDA(cudaSetDevice(device)); }

  // cuda_mpm_model.cu:16-33
  void add_qr_cloth(const T* p, const T* v, size_t nv, const int* t, size_t nt) {
    if (finalized) throw std::invalid_argument("AddQRCloth after Finalize");
    for (size_t i = 0; i < nt * 3; ++i)
      if (t[i] < 0 || (size_t)t[i] >= nv) throw std::invalid_argument("triangle index out of range");
    h_pos.insert(h_pos.end(), p, p + nv * 3);
    h_vel.insert(h_vel.end(), v, v + nv * 3);
    for (size_t i = 0; i < nt * 3; ++i) h_tri.push_back(t[i] + (int)n_verts);
    n_verts += nv;
    n_faces += nt;
    N += nv + nt;
  }

  // Extension: particles that carry their own deformation gradient (F = I at setup) and one of the
  // volumetric materials of volumetric.cuh.  `volume` is the rest volume per particle; the mass is volume * density.
  void add_particles(const T* p, const T* v, const T* volume, size_t n, const gmpm_material* mat) {
    if (finalized) throw std::invalid_argument("AddParticles after Finalize");
    if (!mat) throw std::invalid_argument("AddParticles: null material");
    if (mat->model < GMPM_MAT_FIXED_COROTATED || mat->model > GMPM_MAT_DRUCKER_PRAGER)
      throw std::invalid_argument("AddParticles: unknown material model");
    if (!(mat->youngs > 0) || !(mat->poisson > -1 && mat->poisson < 0.5))
      throw std::invalid_argument("AddParticles: need youngs > 0 and -1 < poisson < 0.5");
    if (n && (!p || !v || !volume)) throw std::invalid_argument("AddParticles: null array");
    size_t m = 0;
    for (; m < materials.size(); ++m)
      if (!std::memcmp(&materials[m], mat, sizeof(gmpm_material))) break;
    if (m == materials.size()) {
      if (m >= (size_t)kMaxMaterials) throw std::invalid_argument("AddParticles: at most 16 distinct materials");
      materials.push_back(*mat);
    }
    h_vpos.insert(h_vpos.end(), p, p + n * 3);
    h_vvel.insert(h_vvel.end(), v, v + n * 3);
    h_vvol.insert(h_vvol.end(), volume, volume + n);
    h_vmat.insert(h_vmat.end(), n, (uint8_t)m);
    n_vol += n;
    N += n;
  }

  // cuda_mpm_model.cu:36-122: every device allocation happens here.
  void finalize() {
    if (finalized) throw std::invalid_argument("Finalize called twice");
    bind_device();
    const size_t F = n_faces, V = n_verts;
    if (N >= (size_t)1 << 31) throw std::invalid_argument("too many particles");
    for (int b = 0; b < 2; ++b) {
      pos[b] = dalloc<T>(3 * N);
      vel[b] = dalloc<T>(3 * N);
      C[b] = dalloc<T>(9 * N);
      vol[b] = dalloc<T>(N);
      pid[b] = dalloc<int>(N);
      pid_of_vis[b] = dalloc<int>(N);
      vis_keys[b] = dalloc<uint32_t>(N);
      vis_ids[b] = dalloc<uint32_t>(N);
      skeys[b] = dalloc<uint32_t>(N);
      svals[b] = dalloc<uint32_t>(N);
      touched_list[b] = dalloc<uint32_t>(k.n_blocks);
    }
    forces = dalloc<T>(3 * N);
    taus = dalloc<T>(6 * N);
    slot_of_pid = dalloc<int>(N);
    vis_of_pid = dalloc<int>(N);
    key = dalloc<uint32_t>(N);
    tri = dalloc<int>(3 * F);
    Fd = dalloc<T>(9 * F);
    Dm_inv = dalloc<T>(4 * F);
    tri_alt = dalloc<int>(3 * F);
    Fd_alt = dalloc<T>(9 * F);
    Dm_inv_alt = dalloc<T>(4 * F);
    fid = dalloc<int>(F);
    fid_alt = dalloc<int>(F);
    face_keys = dalloc<uint32_t>(F);
    sort_scratch = dalloc<uint32_t>(radix_sort_scratch_words(N));
    pb_start = dalloc<uint32_t>(k.n_blocks);
    pb_end = dalloc<uint32_t>(k.n_blocks);
    pb_epoch = dalloc<uint32_t>(k.n_blocks);
    pb_tile = dalloc<uint2>(k.n_blocks);
    pb_list = dalloc<uint32_t>(k.n_blocks);
    // one tile / work record per chunk: every block has at most one partial chunk
    const size_t tile_cap = std::min<size_t>(k.n_blocks, std::max<size_t>(N, 1)) + N / kChunk + 1;
    work_cap = tile_cap;
    pb_work = dalloc<WorkRec>(work_cap);
    tiles = dalloc<T>(tile_cap * 4 * kTileNodes, false);
    g_mass = dalloc<T>(k.n_cells);
    g_vel = dalloc<T>((size_t)k.n_cells * 3);
    g_vstar = dalloc<T>((size_t)k.n_cells * 3);
    touched_epoch = dalloc<uint32_t>(k.n_blocks);
    ctr = dalloc<DevCounters>(1);
    {
      DevCounters init{};
      for (int d = 0; d < 3; ++d) init.win_lo[d] = 0xFFFFFFFFu;
      GMPM_CUDA(cudaMemcpyAsync(ctr, &init, sizeof(init), cudaMemcpyHostToDevice, stream));
      GMPM_CUDA(cudaStreamSynchronize(stream));
    }

    // upload: [faces | verts] (cuda_mpm_model.cu:40-45), planar
    std::vector<T> hp(3 * N, T(0)), hv(3 * N, T(0));
    for (size_t v = 0; v < V; ++v)
      for (int d = 0; d < 3; ++d) {
        hp[d * N + F + v] = h_pos[v * 3 + d];
        hv[d * N + F + v] = h_vel[v * 3 + d];
      }
    std::vector<T> hvol;
    if (n_vol) {  // [faces | verts | volumetric]
      const size_t B0 = F + V;
      hvol.assign(N, T(0));
      for (size_t q = 0; q < n_vol; ++q) {
        for (int d = 0; d < 3; ++d) {
          hp[d * N + B0 + q] = h_vpos[q * 3 + d];
          hv[d * N + B0 + q] = h_vvel[q * 3 + d];
        }
        hvol[B0 + q] = h_vvol[q];
      }
      Fv = dalloc<T>(9 * n_vol, false);
      mat_id = dalloc<uint8_t>(n_vol, false);
      std::vector<T> hF(9 * n_vol, T(0));
      for (size_t q = 0; q < n_vol; ++q) hF[q * 9] = hF[q * 9 + 4] = hF[q * 9 + 8] = T(1);
      GMPM_CUDA(cudaMemcpyAsync(Fv, hF.data(), sizeof(T) * 9 * n_vol, cudaMemcpyHostToDevice, stream));
      GMPM_CUDA(cudaMemcpyAsync(mat_id, h_vmat.data(), n_vol, cudaMemcpyHostToDevice, stream));
      GMPM_CUDA(cudaMemcpyAsync(vol[0], hvol.data(), sizeof(T) * N, cudaMemcpyHostToDevice, stream));
      GMPM_CUDA(cudaStreamSynchronize(stream));  // hF goes out of scope
      for (size_t m = 0; m < materials.size(); ++m) {
        const gmpm_material& g = materials[m];
        MaterialDev<T>& d = mat_table.m[m];
        d.model = g.model;
        d.mu = (T)(g.youngs / (2.0 * (1.0 + g.poisson)));
        d.lambda = (T)(g.youngs * g.poisson / ((1.0 + g.poisson) * (1.0 - 2.0 * g.poisson)));
        const double sphi = std::sin(g.friction_angle_deg * 3.14159265358979323846 / 180.0);
        d.alpha = (T)(std::sqrt(2.0 / 3.0) * 2.0 * sphi / (3.0 - sphi));
        d.cohesion = (T)g.cohesion;
      }
    }
    std::vector<int> ht(3 * F), iota(N);
    for (size_t f = 0; f < F; ++f)
      for (int c = 0; c < 3; ++c) ht[c * F + f] = h_tri[f * 3 + c] + (int)F;
    for (size_t i = 0; i < N; ++i) iota[i] = (int)i;
    GMPM_CUDA(cudaMemcpyAsync(pos[0], hp.data(), sizeof(T) * 3 * N, cudaMemcpyHostToDevice, stream));
    GMPM_CUDA(cudaMemcpyAsync(vel[0], hv.data(), sizeof(T) * 3 * N, cudaMemcpyHostToDevice, stream));
    GMPM_CUDA(cudaMemcpyAsync(tri, ht.data(), sizeof(int) * 3 * F, cudaMemcpyHostToDevice, stream));
    for (int* dst : {pid[0], slot_of_pid, pid_of_vis[0], vis_of_pid})
      GMPM_CUDA(cudaMemcpyAsync(dst, iota.data(), sizeof(int) * N, cudaMemcpyHostToDevice, stream));
    if (F) GMPM_CUDA(cudaMemcpyAsync(fid, iota.data(), sizeof(int) * F, cudaMemcpyHostToDevice, stream));
    if (F) {
      launch("init_fem_state");
      init_fem_state_kernel<T><<<blocks_for(F), kThreads, 0, stream>>>((uint32_t)F, (uint32_t)N, tri, pos[0], vel[0],
                                                                      vol[0], Dm_inv, Fd, k.dx);
      GMPM_CUDA(cudaGetLastError());
    }
    {
      const char* v = std::getenv("GMPM_P2G");  // A/B switch: "tile" (round-1 kernel), "col", "colbulk"
      p2g_variant = !v ? kP2GDefault : !std::strcmp(v, "tile") ? 0 : !std::strcmp(v, "col") ? 1
                                       : !std::strcmp(v, "colbulk") ? 2 : !std::strcmp(v, "pair") ? 3 : -1;
      if (p2g_variant < 0)
        throw std::invalid_argument(std::string("GMPM_P2G=") + v + " is not a P2G variant (tile, col, colbulk, pair)");
      const void* fn = p2g_variant == 0 ? (const void*)p2g_tile_kernel<T>
                       : p2g_variant == 1 ? (const void*)p2g_col_kernel<T, false>
                       : p2g_variant == 2 ? (const void*)p2g_col_kernel<T, true> : (const void*)p2g_pair_kernel<T>;
      p2g_smem = p2g_variant == 0 ? sizeof(P2GSmem<T>) : p2g_variant == 3 ? sizeof(P2GPairSmem<T>) : sizeof(P2GColSmem<T>);
      p2g_threads = p2g_variant == 3 ? kP2GPairThreads : kP2GThreads;
      GMPM_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p2g_smem));
      GMPM_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&p2g_ctas_per_sm, fn, p2g_threads, p2g_smem));
      if (p2g_ctas_per_sm < 1) throw CudaError(cudaErrorLaunchOutOfResources, "the P2G kernel does not fit on an SM");
      if (const char* c = std::getenv("GMPM_P2G_CTAS"))  // development: fewer resident warps than the occupancy allows
        p2g_ctas_per_sm = std::max(1, std::min(p2g_ctas_per_sm, std::atoi(c)));
    }
    GMPM_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&g2p_ctas_per_sm, g2p_kernel<T>, kG2PWarps * 32, 0));
    g2p_ctas_per_sm = std::max(1, std::min(g2p_ctas_per_sm, 8));
    GMPM_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&fem_ctas_per_sm, fem_kernel<T>, kFemThreads, 0));
    fem_ctas_per_sm = std::max(1, fem_ctas_per_sm);
    GMPM_CUDA(cudaStreamSynchronize(stream));  // host staging vectors go out of scope
    finalized = true;
    keys_valid = bin_valid = false;
  }

  void require_finalized() const {
    if (!finalized) throw std::invalid_argument("state is not finalized");
    bind_device();
  }

  // ---- binning ---------------------------------------------------------------------------------
  int key_bits() const { return 3 * k.domain_bits; }

  void ensure_keys() {
    if (keys_valid || N == 0) return;
    join_binning();  // an early binning may still be reading the keys
    launch("compute_keys");
    compute_keys_kernel<T><<<blocks_for(N), kThreads, 0, stream>>>((uint32_t)N, pos[cur], key, k);
    GMPM_CUDA(cudaGetLastError());
    keys_valid = true;
    bin_valid = false;
  }

  const uint32_t* order() const { return svals[sres]; }
  const uint32_t* sorted_keys() const { return skeys[sres]; }

  // binning launched early (calc_fem) on its own stream: see face_keys_kernel
  cudaStream_t bin_stream = nullptr;
  cudaEvent_t ev_prekeys = nullptr, ev_bin_done = nullptr;
  bool bin_async_pending = false;  // the main stream has not yet waited for the early binning
  bool bin_rolled = false;         // ... which has also rolled the touched list on behalf of the coming P2G
  // A/B (profiles/r02_g2p_ab.txt): measured 0.4666 ms/substep with the overlap against 0.4600 without at 1M particles --
  // the cooperative binning and the persistent FEM kernel do not actually share the SMs, and the extra face-key pass
  // costs more than the little that overlaps.  Off unless GMPM_BIN_OVERLAP=1.
  bool bin_overlap = std::getenv("GMPM_BIN_OVERLAP") != nullptr;
  void join_binning() {
    if (!bin_async_pending) return;
    GMPM_CUDA(cudaStreamWaitEvent(stream, ev_bin_done, 0));
    bin_async_pending = false;
  }

  void ensure_binning(bool for_p2g = false, cudaStream_t on = nullptr) {
    ensure_keys();
    prologue_done = false;
    if (N == 0) return;
    if (bin_valid) {
      join_binning();
      if (for_p2g && bin_rolled) prologue_done = true;  // the early binning did p2g_prologue's job
      if (for_p2g) bin_rolled = false;
      return;
    }
    bin_rolled = false;
    if (!on) join_binning();  // an early binning that has been invalidated still owns the sort buffers
    // one cooperative launch: stable sort of (key, slot) on all key bits + block ranges + work records
    ++bin_epoch;
    BinArgs a{};
    a.sort = make_sort_args(key, nullptr, skeys, svals, (uint32_t)N, key_bits(), sort_scratch);
    a.N = (uint32_t)N;
    a.epoch = bin_epoch;
    a.work_cap = (uint32_t)work_cap;
    a.pb_start = pb_start;
    a.pb_end = pb_end;
    a.pb_epoch = pb_epoch;
    a.pb_list = pb_list;
    a.pb_tile = pb_tile;
    a.pb_work = pb_work;
    a.ctr = ctr;
    a.roll_touched = for_p2g ? 1 : 0;
    a.key_bits = key_bits();
    a.window = a.sort.n_passes > 2 ? 1 : 0;
    prologue_done = for_p2g;
    sres = (a.sort.n_passes - 1) & 1;
    if (bin_ctas == 0) bin_ctas = sort_grid_ctas(bin_particles_kernel, sm_count, a.sort.n_tiles);
    void* kargs[] = {&a};
    launch("bin_particles");
    GMPM_CUDA(cudaLaunchCooperativeKernel((void*)bin_particles_kernel, dim3(bin_ctas), dim3(kSortThreads), kargs,
                                          kSortSmemBytes, on ? on : stream));
    g2p_queue_ready = false;
    bin_valid = true;
  }

  // Stable sort of (in_keys, 0..n-1) on `bits` key bits into (ks, vs); returns the result buffer index.
  int sort_iota(const uint32_t* in_keys, uint32_t* ks[2], uint32_t* vs[2], uint32_t n, int bits) {
    SortArgs a = make_sort_args(in_keys, nullptr, ks, vs, n, bits, sort_scratch);
    if (sort_ctas == 0) sort_ctas = sort_grid_ctas(radix_sort_kernel, sm_count, 0xFFFFFFFFu);  // co-resident maximum
    const int ctas = (int)std::min<uint32_t>((uint32_t)sort_ctas, std::max<uint32_t>(a.n_tiles, 64u));
    void* kargs[] = {&a};
    launch("radix_sort");
    GMPM_CUDA(cudaLaunchCooperativeKernel((void*)radix_sort_kernel, dim3(ctas), dim3(kSortThreads), kargs,
                                          kSortSmemBytes, stream));
    return (a.n_passes - 1) & 1;
  }

  // ---- GpuMpmSolver<T>::RebuildMapping (cuda_mpm_solver.cu:18-70) -------------------------------
  // The physical order is maintained by g2p (sorted by full cell key every step); the reference's
  // observable result of a sort -- sorted keys, permutation ids, pids, index_mappings -- is kept as
  // integer bookkeeping of the VISIBLE order, bit-identical to the reference's 16-bit stable sort.
  void rebuild_mapping(bool sort) {
    require_finalized();
    if (N == 0) return;
    ensure_keys();
    if (!sort) {
      vis_keys_valid = false;  // keys = key of each visible particle, ids = iota: materialised lazily
      return;
    }
    launch("gather_visible_keys");
    gather_visible_keys_kernel<<<blocks_for(N), kThreads, 0, stream>>>((uint32_t)N, key, pid_of_vis[vis_cur],
                                                                      slot_of_pid, vis_keys[0], vis_ids[0]);
    GMPM_CUDA(cudaGetLastError());
    const int bits = std::min(key_bits(), 16);  // cuda_mpm_solver.cu:57
    // pass 0 reads vis_keys[0] and must not overwrite it -> it writes the pair [1] first
    uint32_t* ks[2] = {vis_keys[1], vis_keys[0]};
    uint32_t* vs[2] = {vis_ids[1], vis_ids[0]};
    const int r = sort_iota(vis_keys[0], ks, vs, (uint32_t)N, bits);
    vis_res = (r == 0) ? 1 : 0;
    launch("permute_visible");
    permute_visible_kernel<<<blocks_for(N), kThreads, 0, stream>>>((uint32_t)N, vis_ids[vis_res], pid_of_vis[vis_cur],
                                                                  pid_of_vis[vis_cur ^ 1], vis_of_pid);
    GMPM_CUDA(cudaGetLastError());
    vis_cur ^= 1;
    vis_keys_valid = true;
    // pos_stage (SyncParticleStateToCpu -> device-side contact query) is in the visible order that has just been
    // renumbered: gather it again (nothing has moved since the snapshot if the caller follows the driver's order;
    // the HOST mirror keeps the numbering the host saw, as upstream, whose driver never sorts inside this loop)
    if (snapshot_valid) snapshot_positions_device();
  }

  void materialize_visible_keys() {
    if (vis_keys_valid) return;
    ensure_keys();
    launch("gather_visible_keys");
    gather_visible_keys_kernel<<<blocks_for(N), kThreads, 0, stream>>>((uint32_t)N, key, pid_of_vis[vis_cur],
                                                                      slot_of_pid, vis_keys[0], vis_ids[0]);
    GMPM_CUDA(cudaGetLastError());
    vis_res = 0;
    vis_keys_valid = true;
  }

  // Re-orders the face records by the current slot of their face particles (stable sort + gather).
  // Purely a memory-locality measure: no observable result depends on the record order.
  static constexpr int kFaceRefreshEvery = 32;
  void refresh_face_order() {
    const uint32_t F = (uint32_t)n_faces;
    if (F < 2) return;
    launch("face_slot_keys");
    face_slot_keys_kernel<<<blocks_for(F), kThreads, 0, stream>>>(F, fid, slot_of_pid, face_keys);
    GMPM_CUDA(cudaGetLastError());
    int bits = 1;
    while (((size_t)1 << bits) < N) ++bits;
    const int r = sort_iota(face_keys, skeys, svals, F, bits);  // binning scratch: rebuilt after FEM anyway
    bin_valid = false;
    launch("permute_faces");
    permute_faces_kernel<T><<<blocks_for(F), kThreads, 0, stream>>>(F, svals[r], fid, tri, Fd, Dm_inv, fid_alt, tri_alt,
                                                                   Fd_alt, Dm_inv_alt);
    GMPM_CUDA(cudaGetLastError());
    std::swap(fid, fid_alt);
    std::swap(tri, tri_alt);
    std::swap(Fd, Fd_alt);
    std::swap(Dm_inv, Dm_inv_alt);
  }

  // ---- GpuMpmSolver<T>::CalcFemStateAndForce (cuda_mpm_solver.cu:73-84) -------------------------
  void calc_fem(T dt) {
    require_finalized();
    if (N == 0) return;
    ensure_keys();
    if (!forces_zero) {
      note("memset");
      GMPM_CUDA(cudaMemsetAsync(forces, 0, sizeof(T) * 3 * N, stream));
    }
    forces_zero = n_faces == 0;
    if (n_faces) {
      // the particles are in spatial order from the first G2P on: refresh after it, then periodically
      join_binning();  // (a second CalcFemStateAndForce without ParticleToGrid in between)
      if (fem_calls % kFaceRefreshEvery == 1) refresh_face_order();
      ++fem_calls;
      // Binning under the constitutive kernel: the face keys are computed up front, the (cooperative, latency-bound)
      // binning is launched FIRST on its own stream so that all its CTAs are resident, the persistent FEM CTAs fill
      // the rest of the SMs.  Off while profiling (per-kernel times are taken on one stream).
      const bool early = bin_overlap && !prof.on;
      if (early) {
        if (!bin_stream) {
          GMPM_CUDA(cudaStreamCreateWithFlags(&bin_stream, cudaStreamNonBlocking));
          GMPM_CUDA(cudaEventCreateWithFlags(&ev_prekeys, cudaEventDisableTiming));
          GMPM_CUDA(cudaEventCreateWithFlags(&ev_bin_done, cudaEventDisableTiming));
        }
        launch("face_keys");
        face_keys_kernel<T><<<blocks_for(n_faces), kThreads, 0, stream>>>((uint32_t)n_faces, (uint32_t)N, fid, tri, slot_of_pid,
                                                                         pos[cur], key, k);
        GMPM_CUDA(cudaGetLastError());
        GMPM_CUDA(cudaEventRecord(ev_prekeys, stream));
        GMPM_CUDA(cudaStreamWaitEvent(bin_stream, ev_prekeys, 0));
        bin_valid = false;
        ensure_binning(true, bin_stream);
        GMPM_CUDA(cudaEventRecord(ev_bin_done, bin_stream));
        bin_async_pending = true;
        bin_rolled = true;
        prologue_done = false;
      }
      launch("fem");
      const uint32_t ctas = std::min<uint32_t>(blocks_for(n_faces, kFemThreads), persistent_ctas(fem_ctas_per_sm));
      fem_kernel<T><<<ctas, kFemThreads, 0, stream>>>((uint32_t)n_faces, (uint32_t)N, fid, tri, slot_of_pid, vol[cur], C[cur],
                                                      Dm_inv, pos[cur], vel[cur], Fd, forces, taus, early ? nullptr : key, dt, k,
                                                      fem_reverse);
      GMPM_CUDA(cudaGetLastError());
      if (!early) bin_valid = false;  // face particles moved to their centroids
    }
    if (n_vol) {
      launch("volumetric_stress");
      volumetric_stress_kernel<T><<<blocks_for(N, kVolThreads), kVolThreads, 0, stream>>>((uint32_t)N, n_cloth(), pid[cur], C[cur], vol[cur], Fv,
                                                                         mat_id, mat_table, taus, dt);
      GMPM_CUDA(cudaGetLastError());
    }
    fem_valid = true;
  }

  uint32_t persistent_ctas(int ctas_per_sm) const { return (uint32_t)(sm_count * ctas_per_sm); }

  // ---- GpuMpmSolver<T>::ParticleToGrid (cuda_mpm_solver.cu:87-105) ------------------------------
  void p2g(T dt) {
    require_finalized();
    if (N == 0) return;
    ensure_binning(true);
    ++p2g_epoch;
    tl_cur ^= 1;
    if (!prologue_done) {  // binning was still valid: roll the touched list in a launch of its own
      launch("p2g_prologue");
      p2g_prologue_kernel<<<1, 1, 0, stream>>>(ctr);
    }
    prologue_done = false;
    halo_nbr_zeroed_epoch = halo_nbr ? p2g_epoch : 0;  // the tile kernel zeroes the entries it appends
    launch("p2g_tile");
#define GMPM_P2G_ARGS                                                                                            \
  (uint32_t) N, make_uint2((uint32_t)n_faces, n_cloth()), order(), sorted_keys(), pb_work, (uint32_t)work_cap, pos[cur], vel[cur], vol[cur], \
      C[cur], pid[cur], forces, taus, tiles, touched_epoch, touched_list[tl_cur], p2g_epoch, ctr, dt, k, halo_nbr
    if (p2g_variant == 0)
      p2g_tile_kernel<T><<<persistent_ctas(p2g_ctas_per_sm), kP2GThreads, p2g_smem, stream>>>(GMPM_P2G_ARGS);
    else if (p2g_variant == 1)
      p2g_col_kernel<T, false><<<persistent_ctas(p2g_ctas_per_sm), kP2GThreads, p2g_smem, stream>>>(GMPM_P2G_ARGS);
    else if (p2g_variant == 2)
      p2g_col_kernel<T, true><<<persistent_ctas(p2g_ctas_per_sm), kP2GThreads, p2g_smem, stream>>>(GMPM_P2G_ARGS);
    else
      p2g_pair_kernel<T><<<persistent_ctas(p2g_ctas_per_sm), kP2GPairThreads, p2g_smem, stream>>>(GMPM_P2G_ARGS);
#undef GMPM_P2G_ARGS
    g2p_queue_ready = true;  // the prologue reset both work queues
    launch("clean_stale_blocks");
    clean_stale_blocks_kernel<T><<<persistent_ctas(2), kThreads, 0, stream>>>(touched_list[tl_cur ^ 1], ctr,
                                                                             touched_epoch, p2g_epoch, g_mass, g_vel);
    GMPM_CUDA(cudaGetLastError());
    stage = GridStage::kTiles;
  }

  void assemble_momentum() {  // observable P2G result (grid_masses / grid_momentum before UpdateGrid)
    if (halo_recv_pending) halo_recv_now();  // somebody other than UpdateGrid wants the exchanged sums
    if (stage != GridStage::kTiles) return;
    if (halo_nbr) halo_nbr_epoch = halo_nbr_zeroed_epoch == p2g_epoch ? p2g_epoch : 0;
    launch("grid_assemble");
    grid_kernel<T, true, false><<<persistent_ctas(4), kThreads, 0, stream>>>(
        touched_list[tl_cur], ctr, tiles, pb_tile, pb_epoch, bin_epoch, g_mass, g_vel, g_vstar, -1, k,
        halo_nbr_epoch == p2g_epoch ? halo_nbr : nullptr);
    GMPM_CUDA(cudaGetLastError());
    stage = GridStage::kMomentum;
  }

  // ---- GpuMpmSolver<T>::UpdateGrid (cuda_mpm_solver.cu:108-151) ---------------------------------
  void update_grid(int bc) {
    require_finalized();
    if (N == 0) return;
    if (stage == GridStage::kTiles) {
      launch("grid_assemble_update");
      grid_kernel<T, true, true><<<persistent_ctas(4), kThreads, 0, stream>>>(
          touched_list[tl_cur], ctr, tiles, pb_tile, pb_epoch, bin_epoch, g_mass, g_vel, g_vstar, bc, k);
    } else if (halo_recv_pending) {
      halo_recv_pending = false;
      launch("grid_recv_update");
      grid_kernel<T, false, true, 2><<<persistent_ctas(4), kThreads, 0, stream>>>(
          touched_list[tl_cur], ctr, tiles, pb_tile, pb_epoch, bin_epoch, g_mass, g_vel, g_vstar, bc, k, nullptr, nullptr,
          -1, halo_fuse_args());
    } else {
      launch("grid_update");
      grid_kernel<T, false, true><<<persistent_ctas(4), kThreads, 0, stream>>>(
          touched_list[tl_cur], ctr, tiles, pb_tile, pb_epoch, bin_epoch, g_mass, g_vel, g_vstar, bc, k);
    }
    GMPM_CUDA(cudaGetLastError());
    stage = GridStage::kVelocity;
  }

  // ---- GpuMpmSolver<T>::GridToParticle (cuda_mpm_solver.cu:154-161) -----------------------------
  void g2p(T dt) {
    require_finalized();
    if (N == 0) return;
    assemble_momentum();  // only if UpdateGrid was skipped by the caller
    ensure_binning();
    if (!g2p_queue_ready) {
      note("memset");
      GMPM_CUDA(cudaMemsetAsync(&ctr->work_g2p, 0, sizeof(uint32_t), stream));
    }
    g2p_queue_ready = false;
    launch("g2p");
    const int nx = cur ^ 1;
    g2p_kernel<T><<<persistent_ctas(g2p_ctas_per_sm), kG2PWarps * 32, 0, stream>>>(
        (uint32_t)N, order(), pb_work, (uint32_t)work_cap, g_vel, pos[cur], vol[cur], pid[cur], pos[nx], vel[nx], C[nx],
        vol[nx], pid[nx], slot_of_pid, key, forces, ctr, dt, k, nullptr, -1, n_cloth());
    GMPM_CUDA(cudaGetLastError());
    halo_clear_marks();  // an exchange whose parts were started but finished with the plain calls
    forces_zero = true;
    cur = nx;
    keys_valid = true;   // g2p wrote the key of every new slot
    snapshot_valid = false;
    bin_valid = false;
    fem_valid = false;
  }

  // ---- UpdateGrid / GridToParticle in two parts around a shared-block exchange in flight ------------
  // part 0: grid blocks that are NOT exchanged, and the particle chunks whose stencils stay clear of exchanged
  //         blocks: independent of the other ranks, so they run while the all-reduce is in flight;
  // part 1: the rest, after halo_unpack.  Same results as update_grid + g2p (disjoint sets, same kernels).
  void update_grid_part(int bc, int part) {
    require_finalized();
    if (part != 0 && part != 1) throw std::invalid_argument("part must be 0 or 1");
    if (N == 0) return;
    if (!halo_mark) halo_init();
    assemble_momentum();
    if (stage != GridStage::kMomentum) throw std::invalid_argument("update_grid_part must follow ParticleToGrid");
    if (!halo_mark) halo_init();
    if (halo_recv_pending) halo_recv_now();
    halo_set_marks();
    launch(part ? "grid_update_shared" : "grid_update_interior");
    grid_kernel<T, false, true><<<persistent_ctas(4), kThreads, 0, stream>>>(
        touched_list[tl_cur], ctr, tiles, pb_tile, pb_epoch, bin_epoch, g_mass, g_vel, g_vstar, bc, k, nullptr,
        halo_mark, part);
    GMPM_CUDA(cudaGetLastError());
    if (part == 1) stage = GridStage::kVelocity;
  }
  void g2p_part(T dt, int part) {
    require_finalized();
    if (part != 0 && part != 1) throw std::invalid_argument("part must be 0 or 1");
    if (N == 0) return;
    if (!halo_mark) halo_init();
    if (part == 0) {
      if (stage != GridStage::kMomentum) throw std::invalid_argument("g2p_part(0) must follow update_grid_part(0)");
      ensure_binning();
      if (!g2p_queue_ready) {
        note("memset");
        GMPM_CUDA(cudaMemsetAsync(&ctr->work_g2p, 0, 2 * sizeof(uint32_t), stream));  // work_g2p, work_grid
      }
      g2p_queue_ready = false;
    } else if (stage != GridStage::kVelocity) {
      throw std::invalid_argument("g2p_part(1) must follow update_grid_part(1)");
    }
    halo_set_marks();
    launch(part ? "g2p_shared" : "g2p_interior");
    const int nx = cur ^ 1;
    g2p_kernel<T><<<persistent_ctas(g2p_ctas_per_sm), kG2PWarps * 32, 0, stream>>>(
        (uint32_t)N, order(), pb_work, (uint32_t)work_cap, g_vel, pos[cur], vol[cur], pid[cur], pos[nx], vel[nx], C[nx],
        vol[nx], pid[nx], slot_of_pid, key, forces, ctr, dt, k, halo_mark, part, n_cloth());
    GMPM_CUDA(cudaGetLastError());
    if (part == 1) {
      halo_clear_marks();  // the marks of this exchange are spent
      forces_zero = true;
      cur = nx;
      keys_valid = true;
      bin_valid = false;
      fem_valid = false;
      snapshot_valid = false;
    }
  }

  // ---- multi-GPU shared-block exchange (halo.cuh) ---------------------------------------------------
  // plan:      halo_flags -> [all-reduce on halo_stream] -> halo_plan_async      (for the NEXT exchange)
  // exchange:  halo_pack  -> [all-reduce on stream]      -> halo_unpack          (between P2G and UpdateGrid)
  void halo_init() {
    if (halo_flags_d) return;
    for (int i = 0; i < 2; ++i) {
      halo_flags_buf[i] = dalloc<uint8_t>(k.n_blocks);  // zero-initialised
      halo_flags_clean[i] = true;
    }
    halo_flags_d = halo_flags_buf[0];
    halo_own = dalloc<uint8_t>(k.n_blocks);
    halo_nbr = dalloc<uint32_t>(k.n_blocks);
    halo_mark = dalloc<uint8_t>(k.n_blocks);
    halo_cta_counts = dalloc<uint32_t>((k.n_blocks + kHaloBlocksPerCta - 1) / kHaloBlocksPerCta);
    halo_dev_words = dalloc<uint32_t>(8);  // [0..1] counts per plan slot, [2] containment errors, [3] send counter, [4..5] peers per slot
    GMPM_CUDA(cudaHostAlloc((void**)&h_halo_words, 4 * sizeof(uint32_t), cudaHostAllocDefault));
    memset(h_halo_words, 0, 4 * sizeof(uint32_t));
    GMPM_CUDA(cudaStreamCreateWithFlags(&halo_stream, cudaStreamNonBlocking));
    GMPM_CUDA(cudaEventCreateWithFlags(&ev_halo_flags, cudaEventDisableTiming));
    GMPM_CUDA(cudaEventCreateWithFlags(&ev_halo_plan, cudaEventDisableTiming));
    halo_reserve(2048);
  }
  void halo_reserve(size_t blocks) {
    if (blocks <= halo_cap) return;
    size_t cap = std::max<size_t>(2048, halo_cap * 2);
    while (cap < blocks) cap *= 2;
    cap = std::min<size_t>(cap, k.n_blocks);
    GMPM_CUDA(cudaStreamSynchronize(stream));
    GMPM_CUDA(cudaStreamSynchronize(halo_stream));
    for (int i = 0; i < 2; ++i) {
      uint32_t* old = halo_list[i];
      uint8_t* old_sh = halo_sharers[i];
      halo_list[i] = dalloc<uint32_t>(cap, false);
      halo_sharers[i] = dalloc<uint8_t>(cap);
      if (old) {
        GMPM_CUDA(cudaMemcpyAsync(halo_list[i], old, sizeof(uint32_t) * halo_cap, cudaMemcpyDeviceToDevice, stream));
        GMPM_CUDA(cudaMemcpyAsync(halo_sharers[i], old_sh, halo_cap, cudaMemcpyDeviceToDevice, stream));
        GMPM_CUDA(cudaStreamSynchronize(stream));
        dfree(old);
        dfree(old_sh);
      }
    }
    if (halo_packed) dfree(halo_packed);
    halo_packed = dalloc<T>(cap * 256, false);
    halo_cap = cap;
  }
  void require_after_p2g(const char* what) const {
    if (stage != GridStage::kTiles && stage != GridStage::kMomentum)
      throw std::invalid_argument(std::string(what) + " must follow ParticleToGrid and precede UpdateGrid");
  }
  // Flags of the blocks this rank can reach in the next substep (halo.cuh), built on the compute stream
  // from the current touched list (measured: building them on the halo stream as well costs more in
  // contention with FEM / binning than the ~15 us it takes off this stream); the host all-reduces them
  // on the halo stream.
  uint8_t* halo_flags() {
    require_finalized();
    require_after_p2g("halo_flags");
    halo_init();
    // the previous plan chain (halo stream) still reads the flag buffers
    GMPM_CUDA(cudaStreamWaitEvent(stream, ev_halo_plan, 0));
    halo_flags_d = halo_flags_buf[halo_flags_idx];
    if (!halo_flags_clean[halo_flags_idx]) {  // only when halo_flags is repeated without a plan in between
      note("memset");
      GMPM_CUDA(cudaMemsetAsync(halo_flags_d, 0, k.n_blocks, stream));
    }
    halo_flags_clean[halo_flags_idx] = false;
    launch("halo_flags");
    halo_flags_kernel<<<persistent_ctas(2), kThreads, 0, stream>>>(
        touched_list[tl_cur], ctr, (uint32_t)k.domain_size >> 2, halo_own_valid ? halo_own : nullptr,
        (stage == GridStage::kMomentum && halo_nbr_epoch == p2g_epoch) ? halo_nbr : nullptr, halo_flags_d,
        halo_dev_words + 2, (uint8_t)(halo_peer ? 1u << halo_rank : 1u));
    GMPM_CUDA(cudaGetLastError());
    GMPM_CUDA(cudaEventRecord(ev_halo_flags, stream));
    GMPM_CUDA(cudaStreamWaitEvent(halo_stream, ev_halo_flags, 0));
    // this rank's own set, for the containment check of the next halo_flags (which waits for ev_halo_plan): copied
    // on the side stream, before the all-reduce overwrites the flags in place
    GMPM_CUDA(cudaMemcpyAsync(halo_own, halo_flags_d, k.n_blocks, cudaMemcpyDeviceToDevice, halo_stream));
    GMPM_CUDA(cudaEventRecord(ev_halo_plan, halo_stream));  // re-recorded by halo_plan_async; orders a repeated halo_flags
    halo_own_valid = true;
    halo_flags_out = true;
    return halo_flags_d;
  }
  void halo_compact_into(int slot, cudaStream_t st) {
    const uint32_t ctas = (k.n_blocks + kHaloBlocksPerCta - 1) / kHaloBlocksPerCta;
    if (halo_peer)  // forget the list this slot held two exchanges ago (its count is still in the device word)
      halo_index_kernel<<<8, kThreads, 0, st>>>(halo_dev_words + slot, (uint32_t)halo_cap, halo_list[slot],
                                                halo_index[slot], false);
    halo_count_kernel<<<ctas, kHaloThreads, 0, st>>>(k.n_blocks, halo_flags_d, halo_cta_counts, halo_peer,
                                                    halo_dev_words + 4 + slot);
    halo_compact_kernel<<<ctas, kHaloThreads, 0, st>>>(k.n_blocks, halo_flags_d, halo_cta_counts, halo_list[slot],
                                                      (uint32_t)halo_cap, halo_dev_words + slot, halo_peer,
                                                      halo_sharers[slot], halo_dev_words + 4 + slot, halo_rank);
    if (halo_peer)
      halo_index_kernel<<<8, kThreads, 0, st>>>(halo_dev_words + slot, (uint32_t)halo_cap, halo_list[slot],
                                                halo_index[slot], true);
    GMPM_CUDA(cudaGetLastError());
    launches += halo_peer ? 4 : 2;
  }
  // flags now hold the sum over the ranks: build the shared-block list of the next exchange on the side
  // stream; its count lands in pinned memory and is picked up by the next halo_pack.
  void halo_plan_async() {
    require_finalized();
    if (!halo_flags_out) throw std::invalid_argument("halo_plan without halo_flags");
    const int slot = halo_planned ? halo_cur ^ 1 : halo_cur;
    halo_compact_into(slot, halo_stream);
    GMPM_CUDA(cudaMemcpyAsync(h_halo_words + 2 * slot, halo_dev_words + slot, sizeof(uint32_t), cudaMemcpyDeviceToHost,
                              halo_stream));
    GMPM_CUDA(cudaMemcpyAsync(h_halo_words + 2 * slot + 1, halo_dev_words + 2, sizeof(uint32_t),
                              cudaMemcpyDeviceToHost, halo_stream));
    // the other flag buffer belongs to the plan the current exchange has already consumed: zero it for the next
    // halo_flags here, off the compute stream
    if (!halo_flags_clean[halo_flags_idx ^ 1]) {
      GMPM_CUDA(cudaMemsetAsync(halo_flags_buf[halo_flags_idx ^ 1], 0, k.n_blocks, halo_stream));
      halo_flags_clean[halo_flags_idx ^ 1] = true;
    }
    halo_flags_idx ^= 1;
    GMPM_CUDA(cudaEventRecord(ev_halo_plan, halo_stream));
    halo_flags_out = false;
    halo_next_slot = slot;
    halo_plan_pending = true;
  }
  void halo_reset_plan() {
    if (halo_stream) GMPM_CUDA(cudaStreamSynchronize(halo_stream));
    if (stream) GMPM_CUDA(cudaStreamSynchronize(stream));
    halo_planned = halo_plan_pending = halo_flags_out = halo_own_valid = halo_packed_out = false;
    halo_sent = halo_recv_pending = false;
  }
  int halo_next_slot = 0;
  bool halo_plan_pending = false;
  // Start of an exchange: picks up the list planned for it (count through pinned memory), checks containment,
  // assembles the block sums and marks the listed blocks.  Returns the number of shared blocks.
  uint32_t halo_begin_exchange(const char* what, bool assemble = true) {
    require_finalized();
    require_after_p2g(what);
    if (!halo_plan_pending) throw std::invalid_argument(std::string(what) + " without a plan (halo_flags / halo_plan first)");
    GMPM_CUDA(cudaEventSynchronize(ev_halo_plan));  // issued a whole substep ago (or just now at start-up)
    const int slot = halo_next_slot;
    if (h_halo_words[2 * slot + 1] >= kHaloPeerTimeoutCode)
      throw std::runtime_error("halo exchange: a rank waited 20 s for a peer's block values (peer-memory windows): a peer "
                               "stopped, or the ranks do not issue the same sequence of exchanges");
    if (h_halo_words[2 * slot + 1] != 0)
      throw std::runtime_error("halo exchange: a particle block appeared outside the dilated block set planned one "
                               "substep earlier (a particle moved more than one grid block in a substep)");
    uint32_t n = h_halo_words[2 * slot];
    if (halo_peer && n > halo_win_cap)  // the windows are fixed at halo_peer_init: refuse BEFORE touching any list
      throw std::runtime_error("halo exchange: " + std::to_string(n) + " shared blocks, more than the peer windows were "
                               "sized for (" + std::to_string(halo_win_cap) + "): raise cap_blocks in halo_peer_init");
    if (n > halo_cap) {  // rare: the lists were too short; grow and redo the compaction (flags still hold the sums)
      halo_reserve(n);
      halo_compact_into(slot, stream);
      GMPM_CUDA(cudaStreamSynchronize(stream));
    }
    halo_cur = slot;
    halo_planned = true;
    halo_plan_pending = false;
    halo_n = n;
    GMPM_CUDA(cudaStreamWaitEvent(stream, ev_halo_plan, 0));
    if (assemble) assemble_momentum();
    return n;
  }
  // mark[b] = 1 for the blocks of the exchange in flight: only the part-wise UpdateGrid / G2P read them, so they
  // are set when the first part call arrives and cleared after the last
  bool halo_marked = false;
  void halo_set_marks() {
    if (halo_marked || !halo_n) return;
    launch("halo_mark");
    halo_mark_kernel<<<blocks_for(halo_n), kThreads, 0, stream>>>(halo_n, halo_list[halo_cur], 1, halo_mark);
    GMPM_CUDA(cudaGetLastError());
    halo_marked = true;
  }
  void halo_clear_marks() {
    if (!halo_marked) return;
    launch("halo_mark");
    halo_mark_kernel<<<blocks_for(halo_n), kThreads, 0, stream>>>(halo_n, halo_list[halo_cur], 0, halo_mark);
    GMPM_CUDA(cudaGetLastError());
    halo_marked = false;
  }
  // Packs this rank's share of the planned shared blocks; returns their number (known on the host).
  uint32_t halo_pack() {
    if (halo_peer) throw std::invalid_argument("halo_pack: this state exchanges through peer windows (halo_peer_send / recv)");
    const uint32_t n = halo_begin_exchange("halo_pack");
    const int slot = halo_cur;
    if (n) {
      launch("halo_pack");
      halo_pack_kernel<T, false><<<std::min<uint32_t>(n, persistent_ctas(8)), kHaloThreads, 0, stream>>>(
          n, halo_list[slot], touched_epoch, p2g_epoch, g_mass, g_vel, halo_packed);
      GMPM_CUDA(cudaGetLastError());
    }
    halo_packed_out = true;
    return n;
  }
  void halo_unpack() {
    require_finalized();
    if (!halo_packed_out) throw std::invalid_argument("halo_unpack without halo_pack");
    if (halo_n) {
      launch("halo_unpack");
      halo_pack_kernel<T, true><<<std::min<uint32_t>(halo_n, persistent_ctas(8)), kHaloThreads, 0, stream>>>(
          halo_n, halo_list[halo_cur], touched_epoch, p2g_epoch, g_mass, g_vel, halo_packed);
      GMPM_CUDA(cudaGetLastError());
    }
    halo_packed_out = false;
  }

  // ---- peer mode ------------------------------------------------------------------------------------
  // Must precede the first halo_flags.  `cap_blocks` bounds the shared-block list (the window cannot move
  // once the other ranks have mapped it).
  void halo_peer_init(uint32_t rank, uint32_t world, uint32_t cap_blocks) {
    require_finalized();
    if (world < 1 || world > (uint32_t)kHaloMaxPeers || rank >= world)
      throw std::invalid_argument("halo_peer_init: peer mode serves 1..8 ranks (one NVSwitch domain)");
    if (halo_window) throw std::invalid_argument("halo_peer_init: already initialised");
    if (halo_planned || halo_plan_pending || halo_flags_out)
      throw std::invalid_argument("halo_peer_init must precede the first halo_flags");
    halo_init();
    halo_reserve(cap_blocks);
    halo_win_cap = (uint32_t)halo_cap;
    halo_window_bytes = gmpm::halo_window_bytes<T>(world, halo_win_cap);
    GMPM_CUDA(cudaMalloc((void**)&halo_window, halo_window_bytes));
    GMPM_CUDA(cudaMemsetAsync(halo_window, 0, halo_window_bytes, stream));
    GMPM_CUDA(cudaStreamSynchronize(stream));
    for (int i = 0; i < 2; ++i) {
      halo_index[i] = dalloc<uint32_t>(k.n_blocks, false);
      GMPM_CUDA(cudaMemsetAsync(halo_index[i], 0xFF, sizeof(uint32_t) * k.n_blocks, stream));  // kHaloNoIndex
    }
    GMPM_CUDA(cudaStreamSynchronize(stream));
    halo_rank = rank;
    halo_world = world;
    halo_windows.base[rank] = halo_window;
    halo_windows.gpu_scope_fence = std::getenv("GMPM_HALO_GPU_SCOPE_FENCE") != nullptr ? 1u : 0u;
    halo_peer = true;
    halo_own_valid = false;
    halo_xchg = 0;
  }
  void require_peer(const char* what) const {
    require_finalized();
    if (!halo_peer) throw std::invalid_argument(std::string(what) + ": halo_peer_init first");
  }
  void halo_peer_ipc_handle(void* out64) {
    require_peer("halo_peer_ipc_handle");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handle size");
    cudaIpcMemHandle_t h;
    GMPM_CUDA(cudaIpcGetMemHandle(&h, halo_window));
    std::memcpy(out64, &h, sizeof(h));
  }
  // Another rank's window: either its CUDA IPC handle (other process) or its raw device pointer (same process).
  void halo_peer_attach(uint32_t peer, const void* ipc_handle64, void* same_process_ptr) {
    require_peer("halo_peer_attach");
    if (peer >= halo_world || peer == halo_rank) throw std::invalid_argument("halo_peer_attach: bad peer rank");
    if (halo_windows.base[peer]) throw std::invalid_argument("halo_peer_attach: peer already attached");
    if (ipc_handle64) {
      cudaIpcMemHandle_t h;
      std::memcpy(&h, ipc_handle64, sizeof(h));
      void* p = nullptr;
      GMPM_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
      halo_windows.base[peer] = static_cast<unsigned char*>(p);
      halo_window_ipc[peer] = true;
    } else if (same_process_ptr) {
      halo_windows.base[peer] = static_cast<unsigned char*>(same_process_ptr);
    } else {
      throw std::invalid_argument("halo_peer_attach: neither a handle nor a pointer");
    }
  }
  // pack + peer stores + delivery flags, one kernel.  Returns the number of shared blocks.
  uint32_t halo_peer_send() {
    require_peer("halo_peer_send");
    for (uint32_t q = 0; q < halo_world; ++q)
      if (!halo_windows.base[q]) throw std::invalid_argument("halo_peer_send: not every peer window is attached");
    if (halo_sent) throw std::invalid_argument("halo_peer_send: the previous exchange was not received");
    // fused: the grid assembly stores each listed block into its sharers' windows as soon as it is summed
    const bool fuse = halo_fused && stage == GridStage::kTiles;
    const uint32_t n = halo_begin_exchange("halo_peer_send", !fuse);
    if (n > halo_win_cap)
      throw std::runtime_error("halo_peer_send: more shared blocks than the peer windows were sized for (halo_peer_init cap_blocks)");
    ++halo_xchg;
    halo_send_fused = fuse && n;
    if (fuse && !n) assemble_momentum();
    if (halo_send_fused) {
      if (halo_nbr) halo_nbr_epoch = halo_nbr_zeroed_epoch == p2g_epoch ? p2g_epoch : 0;
      launch("grid_assemble_send");
      grid_kernel<T, true, false, 1><<<persistent_ctas(4), kThreads, 0, stream>>>(
          touched_list[tl_cur], ctr, tiles, pb_tile, pb_epoch, bin_epoch, g_mass, g_vel, g_vstar, -1, k,
          halo_nbr_epoch == p2g_epoch ? halo_nbr : nullptr, nullptr, -1, halo_fuse_args());
      GMPM_CUDA(cudaGetLastError());
      stage = GridStage::kMomentum;
    } else if (n) {
      launch("halo_peer_send");
      halo_peer_send_kernel<T><<<n, kHaloPeerThreads, 0, stream>>>(
          n, halo_list[halo_cur], halo_sharers[halo_cur], touched_epoch, p2g_epoch, g_mass, g_vel, halo_windows, halo_rank,
          halo_world, halo_win_cap, halo_xchg, halo_dev_words + 4 + halo_cur, halo_dev_words + 3);
      GMPM_CUDA(cudaGetLastError());
    }
    halo_sent = true;
    return n;
  }
  // waits for the sharers' deliveries on the device and writes the rank-ordered sums into this rank's blocks
  void halo_peer_recv() {
    require_peer("halo_peer_recv");
    if (!halo_sent) throw std::invalid_argument("halo_peer_recv without halo_peer_send");
    halo_sent = false;
    if (halo_n && halo_fused) {  // the next UpdateGrid waits for the sharers block by block and adds their values
      halo_recv_pending = true;
      return;
    }
    halo_recv_now();
  }
  bool halo_send_fused = false;
  HaloFuse<T> halo_fuse_args() const {
    HaloFuse<T> hf{};
    hf.index = halo_index[halo_cur];
    hf.list = halo_list[halo_cur];
    hf.n_list = halo_n;
    hf.touched_epoch = touched_epoch;
    hf.p2g_epoch = p2g_epoch;
    hf.sharers = halo_sharers[halo_cur];
    hf.win = halo_windows;
    hf.rank = halo_rank;
    hf.world = halo_world;
    hf.cap = halo_win_cap;
    hf.xchg = halo_xchg;
    hf.peers = halo_dev_words + 4 + halo_cur;
    hf.done = halo_dev_words + 3;
    hf.err = halo_dev_words + 2;
    return hf;
  }
  // the receive as a kernel of its own (unfused mode, or a consumer other than UpdateGrid needs the sums)
  void halo_recv_now() {
    halo_recv_pending = false;
    if (halo_n) {
      launch("halo_peer_recv");
      halo_peer_recv_kernel<T><<<halo_n, kHaloPeerThreads, 0, stream>>>(
          halo_n, halo_list[halo_cur], halo_sharers[halo_cur], touched_epoch, p2g_epoch, g_mass, g_vel, halo_window, halo_rank,
          halo_world, halo_win_cap, halo_xchg, halo_dev_words + 2);
      GMPM_CUDA(cudaGetLastError());
    }
  }

  cudaEvent_t timers[16]{};
  void timer_mark(int slot) {
    if (slot < 0 || slot >= 16) throw std::invalid_argument("timer slot out of range");
    bind_device();
    if (!timers[slot]) GMPM_CUDA(cudaEventCreate(&timers[slot]));
    GMPM_CUDA(cudaEventRecord(timers[slot], stream));
  }
  double timer_elapsed_ms(int a, int b) {
    if (a < 0 || a >= 16 || b < 0 || b >= 16 || !timers[a] || !timers[b]) throw std::invalid_argument("timer slot not marked");
    bind_device();
    GMPM_CUDA(cudaEventSynchronize(timers[b]));
    float ms = 0.f;
    GMPM_CUDA(cudaEventElapsedTime(&ms, timers[a], timers[b]));
    return (double)ms;
  }

  void sync() {
    bind_device();
    if (bin_async_pending) GMPM_CUDA(cudaStreamSynchronize(bin_stream));
    GMPM_CUDA(cudaStreamSynchronize(stream));
    if (prof.on) prof.flush(stream);
  }

  void run_substeps(int n, T dt, int bc, int sort_every) {
    for (int i = 0; i < n; ++i) {
      rebuild_mapping(sort_every > 0 && (i % sort_every) == 0);
      calc_fem(dt);
      p2g(dt);
      update_grid(bc);
      g2p(dt);
    }
  }

  // ---- host I/O --------------------------------------------------------------------------------
  void* io(size_t bytes) {
    if (bytes > io_cap) {
      GMPM_CUDA(cudaStreamSynchronize(stream));
      if (io_buf) cudaFree(io_buf);
      io_buf = nullptr;
      io_cap = 0;
      GMPM_CUDA(cudaMalloc(&io_buf, bytes));
      io_cap = bytes;
    }
    return io_buf;
  }
  void to_host(void* dst, const void* src, size_t bytes) {
    GMPM_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, stream));
    GMPM_CUDA(cudaStreamSynchronize(stream));
  }

  // GpuMpmSolver<T>::SyncParticleStateToCpu (cuda_mpm_solver.cu:187-191)
  void positions_to_host(T* out) {
    require_finalized();
    if (N == 0) return;
    T* buf = static_cast<T*>(io(sizeof(T) * 3 * N));
    launch("gather_positions");
    gather_particle_field_kernel<T><<<blocks_for(N), kThreads, 0, stream>>>((uint32_t)N, 3, pos[cur],
                                                                           pid_of_vis[vis_cur], slot_of_pid, buf);
    GMPM_CUDA(cudaGetLastError());
    to_host(out, buf, sizeof(T) * 3 * N);
    if (prof.on) prof.flush(stream);
  }

  // Asynchronous form of SyncParticleStateToCpu.  The driver needs positions_host() only when it builds
  // the contact pairs, four solver calls later (deformable_driver.h:243 vs :247), so the PCIe copy can
  // run under FEM / P2G / UpdateGrid: the gather to the reference's layout and visible order runs in
  // order on the compute stream (positions as of THIS call), the D2H runs on a copy stream into a
  // library-owned pinned mirror, and wait_positions() blocks on that copy alone.
  // The device half alone: positions of this moment in the reference's layout and visible order in pos_stage
  // (what the device-side contact query reads when nobody asked for a host copy).
  void snapshot_positions_device() {
    require_finalized();
    if (N == 0) return;
    if (!pos_stage) pos_stage = dalloc<T>(3 * N, false);
    if (copy_pending) GMPM_CUDA(cudaStreamWaitEvent(stream, ev_copy, 0));  // staging buffer still being read
    launch("gather_positions");
    gather_particle_field_kernel<T><<<blocks_for(N), kThreads, 0, stream>>>((uint32_t)N, 3, pos[cur],
                                                                           pid_of_vis[vis_cur], slot_of_pid, pos_stage);
    GMPM_CUDA(cudaGetLastError());
    snapshot_valid = true;
  }
  void ensure_copy_stream() {
    if (copy_stream) return;
    GMPM_CUDA(cudaStreamCreateWithFlags(&copy_stream, cudaStreamNonBlocking));
    GMPM_CUDA(cudaEventCreateWithFlags(&ev_gather, cudaEventDisableTiming));
    GMPM_CUDA(cudaEventCreateWithFlags(&ev_copy, cudaEventDisableTiming));
    GMPM_CUDA(cudaEventCreateWithFlags(&ev_count, cudaEventDisableTiming));
    GMPM_CUDA(cudaHostAlloc((void**)&h_pos_mirror, sizeof(T) * 3 * N, cudaHostAllocDefault));
  }
  static constexpr double kFarAway = 1e30;  // positions_host() of a particle outside every declared query bound
  // Declares the bounds of the rigid geometry the host contact query tests against (n axis-aligned boxes, lo/hi
  // interleaved: 6 scalars each, inflated by the caller as it sees fit).  From now on the snapshot moves only the
  // particles inside one of them to the host; positions_host() of every other particle reads kFarAway, for which any
  // signed-distance query is positive, so the host loop (deformable_driver.h:138-190) builds the pairs it would have
  // built from the full snapshot.  n = 0: nothing but a count crosses PCIe.
  void set_query_bounds(const T* lo_hi, size_t n) {
    require_finalized();
    if (n > (size_t)kMaxQueryBounds) throw std::invalid_argument("set_query_bounds: at most 64 boxes");
    if (n && !lo_hi) throw std::invalid_argument("set_query_bounds: null bounds");
    if (N == 0) return;
    ensure_copy_stream();
    wait_positions();
    if (!d_qb) {
      d_qb = dalloc<QueryBounds<T>>(1);
      cand_count = dalloc<uint32_t>(1);
      cand_idx = dalloc<uint32_t>(N, false);
      cand_pos = dalloc<T>(3 * N, false);
      GMPM_CUDA(cudaHostAlloc((void**)&h_cand_count, sizeof(uint32_t), cudaHostAllocDefault));
      GMPM_CUDA(cudaHostAlloc((void**)&h_cand_idx, sizeof(uint32_t) * N, cudaHostAllocDefault));
      GMPM_CUDA(cudaHostAlloc((void**)&h_cand_pos, sizeof(T) * 3 * N, cudaHostAllocDefault));
    }
    QueryBounds<T> h{};
    h.n = (int)n;
    for (size_t b = 0; b < n; ++b)
      for (int c = 0; c < 3; ++c) {
        h.lo[b][c] = lo_hi[b * 6 + c];
        h.hi[b][c] = lo_hi[b * 6 + 3 + c];
      }
    GMPM_CUDA(cudaMemcpyAsync(d_qb, &h, sizeof(h), cudaMemcpyHostToDevice, stream));
    GMPM_CUDA(cudaStreamSynchronize(stream));  // h is a stack object
    if (!cull_on) {
      for (size_t i = 0; i < 3 * N; ++i) h_pos_mirror[i] = T(kFarAway);
      h_prev_idx.clear();
    }
    cull_on = true;
    cand_last = 0;  // new bounds: nothing to speculate on
  }
  void clear_query_bounds() {
    if (cull_on) wait_positions();
    cull_on = false;
  }
  void positions_async() {
    require_finalized();
    if (N == 0) return;
    ensure_copy_stream();
    if (cull_on) {
      if (cull_pending) wait_positions();  // the previous candidate list was never consumed
      if (!pos_stage) pos_stage = dalloc<T>(3 * N, false);
      if (copy_pending) GMPM_CUDA(cudaStreamWaitEvent(stream, ev_copy, 0));
      GMPM_CUDA(cudaMemsetAsync(cand_count, 0, sizeof(uint32_t), stream));
      launch("cull_positions");
      cull_positions_kernel<T><<<blocks_for(N), kThreads, 0, stream>>>((uint32_t)N, pos[cur], pid_of_vis[vis_cur], slot_of_pid,
                                                                      pos_stage, d_qb, cand_count, cand_idx, cand_pos);
      GMPM_CUDA(cudaGetLastError());
      GMPM_CUDA(cudaEventRecord(ev_gather, stream));
      GMPM_CUDA(cudaStreamWaitEvent(copy_stream, ev_gather, 0));
      GMPM_CUDA(cudaMemcpyAsync(h_cand_count, cand_count, sizeof(uint32_t), cudaMemcpyDeviceToHost, copy_stream));
      GMPM_CUDA(cudaEventRecord(ev_count, copy_stream));
      // speculative fetch: the candidate set changes slowly, so as many candidates as last time (+ 1/8) are copied
      // right away, under the solver calls that follow; wait_positions() fetches the remainder, if any
      cand_fetched = (uint32_t)std::min<size_t>(N, (size_t)cand_last + cand_last / 8 + 64);
      if (cand_last == 0) cand_fetched = 0;
      if (cand_fetched) {
        GMPM_CUDA(cudaMemcpyAsync(h_cand_idx, cand_idx, sizeof(uint32_t) * cand_fetched, cudaMemcpyDeviceToHost, copy_stream));
        GMPM_CUDA(cudaMemcpyAsync(h_cand_pos, cand_pos, sizeof(T) * 3 * cand_fetched, cudaMemcpyDeviceToHost, copy_stream));
      }
      GMPM_CUDA(cudaEventRecord(ev_copy, copy_stream));
      copy_pending = true;
      cull_pending = true;
      snapshot_valid = true;
      return;
    }
    snapshot_positions_device();
    GMPM_CUDA(cudaEventRecord(ev_gather, stream));
    GMPM_CUDA(cudaStreamWaitEvent(copy_stream, ev_gather, 0));
    GMPM_CUDA(cudaMemcpyAsync(h_pos_mirror, pos_stage, sizeof(T) * 3 * N, cudaMemcpyDeviceToHost, copy_stream));
    GMPM_CUDA(cudaEventRecord(ev_copy, copy_stream));
    copy_pending = true;
    snapshot_valid = true;
    snapshot_d2h_bytes = sizeof(T) * 3 * N;
  }
  // Blocks until the last positions_async() copy has landed; returns the pinned mirror (3N, visible order).
  const T* wait_positions() {
    require_finalized();
    if (cull_pending) {
      // the count is long there (the driver asks four solver calls later): fetch exactly that many candidates
      GMPM_CUDA(cudaEventSynchronize(ev_count));
      const uint32_t n = *h_cand_count;
      if (n > cand_fetched) {
        const uint32_t f = cand_fetched, r = n - cand_fetched;
        GMPM_CUDA(cudaMemcpyAsync(h_cand_idx + f, cand_idx + f, sizeof(uint32_t) * r, cudaMemcpyDeviceToHost, copy_stream));
        GMPM_CUDA(cudaMemcpyAsync(h_cand_pos + 3 * (size_t)f, cand_pos + 3 * (size_t)f, sizeof(T) * 3 * r,
                                  cudaMemcpyDeviceToHost, copy_stream));
        GMPM_CUDA(cudaEventRecord(ev_copy, copy_stream));
      }
      GMPM_CUDA(cudaEventSynchronize(ev_copy));
      cand_last = n;
      for (uint32_t i : h_prev_idx)
        h_pos_mirror[3 * (size_t)i] = h_pos_mirror[3 * (size_t)i + 1] = h_pos_mirror[3 * (size_t)i + 2] = T(kFarAway);
      h_prev_idx.assign(h_cand_idx, h_cand_idx + n);
      for (uint32_t j = 0; j < n; ++j) {
        const size_t i = h_cand_idx[j];
        h_pos_mirror[3 * i] = h_cand_pos[3 * (size_t)j];
        h_pos_mirror[3 * i + 1] = h_cand_pos[3 * (size_t)j + 1];
        h_pos_mirror[3 * i + 2] = h_cand_pos[3 * (size_t)j + 2];
      }
      snapshot_d2h_bytes = sizeof(uint32_t) + (uint64_t)std::max(n, cand_fetched) * (sizeof(uint32_t) + 3 * sizeof(T));
      cull_pending = false;
      copy_pending = false;
      return h_pos_mirror;
    }
    if (copy_pending) {
      GMPM_CUDA(cudaEventSynchronize(ev_copy));
      copy_pending = false;
    }
    return h_pos_mirror;
  }

  // GpuMpmState<T>::DumpCpuState (cuda_mpm_model.cu:245-265)
  void dump_cpu_state(T* vert_pos, int* tri_out) {
    require_finalized();
    if (n_verts) {
      T* buf = static_cast<T*>(io(sizeof(T) * 3 * n_verts));
      launch("gather_vertex_positions");
      gather_vertex_positions_kernel<T><<<blocks_for(n_verts), kThreads, 0, stream>>>(
          (uint32_t)n_verts, (uint32_t)n_faces, (uint32_t)N, pos[cur], slot_of_pid, buf);
      GMPM_CUDA(cudaGetLastError());
      to_host(vert_pos, buf, sizeof(T) * 3 * n_verts);
    }
    std::memcpy(tri_out, h_tri.data(), sizeof(int) * 3 * n_faces);
  }

  uint32_t touched_count_host() {
    require_finalized();
    uint32_t c = 0;
    to_host(&c, &ctr->n_touched, sizeof(uint32_t));
    return c;
  }
  uint32_t pblock_count_host() {
    require_finalized();
    uint32_t c = 0;
    to_host(&c, &ctr->n_pblocks, sizeof(uint32_t));
    return c;
  }

  size_t read(const char* name, void* out, size_t cap);  // getters.cuh
  void run_driver_substeps(int n, T dt, int bc, int sort_every, const gmpm_analytic_geometry* geoms, size_t n_geoms,
                           int query_every, T mu, T stiffness, T damping, bool exact, int max_it,
                           gmpm_driver_stats* out);      // contact_gen.cuh
  void invalidate_after_particle_edit();                 // migrate.cuh
  size_t migrate_out(int axis, T lo, T hi, const T* park, T* records, int* pids_out, size_t cap);
  void migrate_in(const T* records, size_t n, int* pids_out);
  void migrate_count(int axis, T lo, T hi, const double* edges, size_t n_edges, size_t* leavers, size_t* n_free);
  size_t checkpoint_bytes() const;                       // checkpoint.cuh
  size_t save_state(void* out, size_t cap);
  void load_state(const void* in, size_t bytes);
  // contact.cuh
  void copy_contact_pairs(size_t n, const uint32_t* mpm_id, const uint32_t* rigid_id, const T* dist, const T* normal,
                          const T* cpos, const T* rigid_v, const T* rigid_p);
  void update_contact(T dt, T mu, T stiffness, T damping, bool exact, int max_it, gmpm_contact_stats* stats);
  void reallocate_external_bodies(size_t n);
  void external_body_force_to_host(T* tau, T* f);
  void reallocate_contacts(size_t n);
  size_t generate_contact_pairs(const gmpm_analytic_geometry* geoms, size_t n_geoms);
  T* contact_grid_dense(const T* src, int comps);
  void dfree(void* p);
};

}  // namespace gmpm
