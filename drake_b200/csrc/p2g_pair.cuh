// p2g_pair.cuh -- P2G tile kernel with TWO warps per tile (replaces particle_to_grid_kernel :418-543 and
// clean_grid_kernel :585-602; same contract and the same arithmetic as p2g_tile_kernel in kernels.cuh).
//
// Why: the round-2 A/B (profiles/r02_p2g_ab.txt) showed that neither P2G kernel is bound by a pipe: with one warp per
// 6^3 x 4 tile + stage (19.7 KB of shared memory at fp64) an SM holds 11 warps, every one of them spends ~7 cycles per
// instruction in dependent-issue and scoreboard stalls, and the issue slots are 34-38 % used.  The way to more warps
// without more shared memory is to let two warps share ONE tile and ONE stage:
//   phase (a)  each warp derives HALF of the batch's 32 particles (rows 16 w .. 16 w + 15) into the common stage;
//   phase (b)  lanes = the 27 stencil nodes in both warps, but warp 0 accumulates channels (mass, p_x) and warp 1
//              (p_y, p_z): disjoint tile addresses, so the run flushes need no ordering between the warps, and each warp
//              loads 32 (warp 0) / 48 (warp 1) operand bytes per particle instead of 64;
//   two CTA barriers per batch (stage full / stage free) and one per item (flushes done -> tile store).
// The work-queue ticket is taken by warp 0 and handed over through shared memory.  Per-channel arithmetic and summation
// order are exactly those of p2g_tile_kernel, so the results are bit-identical to it.
#pragma once
#include "kernels.cuh"

namespace gmpm {

constexpr int kP2GPairThreads = 64;

template <typename T>
struct P2GPairSmem {
  T tile[4][kTileNodes];
  T stage[32][kStageStride<T>];
  uint32_t ticket[2];
};

template <typename T>
__global__ void __launch_bounds__(kP2GPairThreads, 8)
p2g_pair_kernel(uint32_t N, uint2 n_faces, const uint32_t* __restrict__ order,
                const uint32_t* __restrict__ sorted_key, const WorkRec* __restrict__ pb_work, uint32_t work_cap,
                const T* __restrict__ pos, const T* __restrict__ vel, const T* __restrict__ vol,
                const T* __restrict__ C, const int* __restrict__ pid, const T* __restrict__ forces,
                const T* __restrict__ taus, T* __restrict__ tiles, uint32_t* __restrict__ touched_epoch,
                uint32_t* __restrict__ touched_list, uint32_t epoch, DevCounters* ctr, T dt, Consts<T> k,
                uint32_t* __restrict__ nbr_list = nullptr) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  P2GPairSmem<T>& sm = *reinterpret_cast<P2GPairSmem<T>*>(smem_raw);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  T* tile = &sm.tile[0][0];
  const uint32_t n_big = ctr->n_big, n_work = n_big + ctr->n_small;
  if (n_work == 0) return;
  const int ni = lane / 9, nj = (lane / 3) % 3, nk = lane % 3;  // node of this lane (lane < 27)
  const bool node_lane = lane < 27;
  const int ij = node_lane ? ni * 3 + nj : 0, kz = node_lane ? nk : 0;
  const uint32_t blocks_per_axis = (uint32_t)k.domain_size >> 2;
  const bool deriver = lane < 16;              // phase (a): lanes 0-15 of warp w own rows 16 w + lane
  const int my_row = warp * 16 + (lane & 15);

  constexpr uint32_t kNone = 0xFFFFFFFFu;
  using Item = WorkRec;
  int tslot = 0;
  auto ticket = [&]() -> uint32_t {
    uint32_t t = 0;
    if (threadIdx.x == 0) t = atomicAdd(&ctr->work_p2g, 1u);
    return t;
  };
  // both warps call this at the same points of the (identical) control flow
  auto resolve = [&](uint32_t t_thread0) -> Item {
    if (threadIdx.x == 0) sm.ticket[tslot] = t_thread0;
    __syncthreads();
    const uint32_t t = sm.ticket[tslot];
    tslot ^= 1;  // the slot is rewritten two hand-overs later: at least one barrier in between
    if (t >= n_work) return Item{kNone, 0u, 0u, 0u};
    return pb_work[t < n_big ? t : work_cap - 1u - (t - n_big)];
  };
  Item cur = resolve(ticket());
  if (cur.tile == kNone) return;
  Item nxt = resolve(ticket());
  // per lane: the cell key of batch row `lane` (run heads; all 32 rows, both warps) and, for the deriving lanes, the
  // slot of the particle they own -- indices two batches ahead, the record one batch ahead (as p2g_tile_kernel)
  auto batch_idx = [&](const Item& c, uint32_t p0, uint32_t& s, uint32_t& key) -> void {
    s = 0; key = 0xFFFFFFFFu;
    if (c.tile == kNone) return;
    if (p0 + lane < c.end) key = sorted_key[p0 + lane];
    if (deriver && p0 + my_row < c.end) s = order[p0 + my_row];
  };
  auto next_pos = [&](const Item& c, const Item& n, uint32_t p0, const Item*& oc, uint32_t& op) -> void {
    if (p0 + 32 < c.end) { oc = &c; op = p0 + 32; }
    else { oc = &n; op = n.begin; }
  };
  uint32_t s_a, key_a, s_b, key_b;
  batch_idx(cur, cur.begin, s_a, key_a);
  const Item* c1; uint32_t p1;
  next_pos(cur, nxt, cur.begin, c1, p1);
  batch_idx(*c1, p1, s_b, key_b);
  P2GRaw<T> raw;
  bool v_raw = deriver && cur.begin + my_row < cur.end;
  p2g_load(raw, v_raw, s_a, N, n_faces, pos, vel, vol, C, pid, forces, taus);
  uint32_t key_cur = key_a;  // cell key of batch row `lane` of the batch being processed

  for (;;) {
    const uint32_t blk = cur.block;
    for (int i = threadIdx.x; i < 4 * kTileNodes; i += kP2GPairThreads) tile[i] = T(0);
    uint32_t reach = 0;
    T acc0 = T(0), acc1 = T(0);   // this warp's two channels
    int run_cell = -1, last_cell = -1;
    const uint32_t t_pending = ticket();  // the item after `nxt`
    Item nn{kNone, 0u, 0u, 0u};

    for (uint32_t p0 = cur.begin; p0 < cur.end; p0 += 32) {
      const bool last_batch = p0 + 32 >= cur.end;
      // ---- phase (a): lanes 0-15 of each warp derive their particle's row
      const bool valid = p0 + lane < cur.end;
      const int my_cell = (int)(key_cur & 63u);
      if (v_raw) reach |= p2g_derive(raw, &sm.stage[my_row][0], dt, k);
      int prev = __shfl_up_sync(kFullMask, my_cell, 1);
      if (lane == 0) prev = last_cell;
      const uint32_t heads = __ballot_sync(kFullMask, valid && my_cell != prev);
      const int cnt = (int)min(32u, cur.end - p0);
      last_cell = __shfl_sync(kFullMask, my_cell, cnt - 1);
      // ---- prefetch: records of the next batch, indices of the one after
      if (last_batch && nxt.tile != kNone) nn = resolve(t_pending);  // (contains a CTA barrier: also "stage full")
      else __syncthreads();                                          // stage full
      {
        const Item* cb; uint32_t pb;
        next_pos(cur, nxt, p0, cb, pb);                              // the batch whose indices are in (s_b, key_b)
        v_raw = deriver && cb->tile != kNone && pb + my_row < cb->end;
        p2g_load(raw, v_raw, s_b, N, n_faces, pos, vel, vol, C, pid, forces, taus);
        key_cur = key_b;
        const Item* cc; uint32_t pc;
        if (cb->tile == kNone) { s_b = 0; key_b = 0xFFFFFFFFu; }
        else {
          next_pos(*cb, cb == &cur ? nxt : nn, pb, cc, pc);
          batch_idx(*cc, pc, s_b, key_b);
        }
      }
      // ---- phase (b): lanes = stencil nodes; warp 0: (mass, p_x), warp 1: (p_y, p_z)
      auto row_of = [&](int q) { return &sm.stage[q][0]; };
      // operands of particle q for this warp: A' and z always; (G0, zb0) or (G1, G2, zb1, zb2)
      T aA, aG0, aG1, zZ, zB0, zB1;
      auto load_ops = [&](int q, T& oA, T& oG0, T& oG1, T& oZ, T& oB0, T& oB1) {
        const T* r = row_of(q);
        const T* a = r + ij * 4;
        const T* z = r + (9 + kz) * 4;
        if (warp == 0) {
          const typename Pair<T>::type pa = *reinterpret_cast<const typename Pair<T>::type*>(a);      // A', G0
          const typename Pair<T>::type pz = *reinterpret_cast<const typename Pair<T>::type*>(z);      // z, zb0
          oA = pa.x; oG0 = pa.y; oG1 = T(0); oZ = pz.x; oB0 = pz.y; oB1 = T(0);
        } else {
          const typename Pair<T>::type pg = *reinterpret_cast<const typename Pair<T>::type*>(a + 2);  // G1, G2
          const typename Pair<T>::type pb2 = *reinterpret_cast<const typename Pair<T>::type*>(z + 2); // zb1, zb2
          oA = a[0]; oG0 = pg.x; oG1 = pg.y; oZ = z[0]; oB0 = pb2.x; oB1 = pb2.y;
        }
      };
      load_ops(0, aA, aG0, aG1, zZ, zB0, zB1);
      for (int q = 0; q < cnt; ++q) {
        const int qn = q + 1 < cnt ? q + 1 : q;
        T nA, nG0, nG1, nZ, nB0, nB1;
        load_ops(qn, nA, nG0, nG1, nZ, nB0, nB1);
        if ((heads >> q) & 1u) {
          if (run_cell >= 0 && node_lane) {
            const int n = (((run_cell >> 4) & 3) + ni) * 36 + (((run_cell >> 2) & 3) + nj) * 6 + ((run_cell & 3) + nk);
            tile[(2 * warp) * kTileNodes + n] += acc0;
            tile[(2 * warp + 1) * kTileNodes + n] += acc1;
          }
          __syncwarp();
          acc0 = acc1 = T(0);
          run_cell = __shfl_sync(kFullMask, my_cell, q);
        }
        if (warp == 0) {
          acc0 += aA * zZ;                  // mass
          acc1 += aG0 * zZ + aA * zB0;      // p_x
        } else {
          acc0 += aG0 * zZ + aA * zB0;      // p_y  (aG0 = G1, zB0 = zb1)
          acc1 += aG1 * zZ + aA * zB1;      // p_z
        }
        aA = nA; aG0 = nG0; aG1 = nG1; zZ = nZ; zB0 = nB0; zB1 = nB1;
      }
      __syncthreads();  // stage free
    }
    if (run_cell >= 0 && node_lane) {
      const int n = (((run_cell >> 4) & 3) + ni) * 36 + (((run_cell >> 2) & 3) + nj) * 6 + ((run_cell & 3) + nk);
      tile[(2 * warp) * kTileNodes + n] += acc0;
      tile[(2 * warp + 1) * kTileNodes + n] += acc1;
    }
    __syncthreads();  // both warps' flushes are in the tile
    // ---- store the tile (coalesced) and publish the touched grid blocks (each warp those its particles reached)
    T* out = tiles + (size_t)cur.tile * (4 * kTileNodes);
    for (int i = threadIdx.x; i < 4 * kTileNodes; i += kP2GPairThreads) out[i] = tile[i];
    reach = __reduce_or_sync(kFullMask, reach);
    if (lane < 8 && ((reach >> lane) & 1u)) {
      uint32_t bx, by, bz;
      block_coords(blk, bx, by, bz);
      bx += (lane >> 2) & 1u;
      by += (lane >> 1) & 1u;
      bz += lane & 1u;
      if (bx < blocks_per_axis && by < blocks_per_axis && bz < blocks_per_axis) {
        const uint32_t g = morton3(bx, by, bz);
        if (atomicExch(&touched_epoch[g], epoch) != epoch) {
          const uint32_t idx = atomicAdd(&ctr->n_touched, 1u);
          touched_list[idx] = g;
          if (nbr_list) nbr_list[idx] = 0u;  // filled by the grid assembly (multi-GPU planning)
        }
      }
    }
    if (nxt.tile == kNone) break;
    cur = nxt;
    nxt = nn;
  }
}

}  // namespace gmpm
