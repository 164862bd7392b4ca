// C ABI of the prophet_b200 device layer (include/prophet_b200.h): context, set-up, training epoch, MD step.
#include "../../include/prophet_b200.h"
#include "pp_common.cuh"
#include "pp_kernels.h"
#include "pp_comm.h"
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <map>

// ---------------------------------------------------------------------------------------------------------
static thread_local char g_err[1024] = "";
void pp_set_error(const char* fmt, ...)
{
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
extern "C" const char* pp_last_error(void) { return g_err; }
extern "C" int pp_version(void) { return 100; }

// growable device buffer
template <typename T>
struct DevBuf {
  T* p = nullptr;
  size_t cap = 0;
  int ensure(size_t n, size_t keep, cudaStream_t st)
  {
    if (n <= cap) return 0;
    size_t ncap = std::max(n, cap + cap / 2);
    T* q = nullptr;
    PP_CUDA(cudaMalloc(&q, ncap * sizeof(T)));
    if (keep && p) PP_CUDA(cudaMemcpyAsync(q, p, keep * sizeof(T), cudaMemcpyDeviceToDevice, st));
    if (p) {
      PP_CUDA(cudaStreamSynchronize(st));
      PP_CUDA(cudaFree(p));
    }
    p = q;
    cap = ncap;
    return 0;
  }
  void release()
  {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
};

struct NetHost {
  bool set = false;
  int n_layers = 0, P = 0, p_off = 0, n_nodes = 0;
  std::vector<int> sizes, tf;
  std::vector<double> params, mean, scale;
  DevBuf<double> d_mean, d_scale;
  DevBuf<int> d_tf;
  DevBuf<unsigned char> d_dw_tiles, d_dw_tiles2, d_dw_tiles3, d_dw_tiles32;
  bool has_tiles32 = false;            // eligible for the FP32-mode training kernel (pp_mlp32.cu)
  bool has_tiles2 = false;
  bool has_tiles3 = false;             // eligible for the two-phase kernels with stored activations (pp_mlpw.cu)
  int tile_atoms = 0;
  NetDev dev;
};

struct pp_context {
  int device = 0;
  int fp_mode = 64;                    // 32: training epochs of eligible systems run in single precision (pp_mlp32.cu)
  DevBuf<float> d_G32;                 // FP32 mode: normalised descriptors [atoms][NGp]
  DevBuf<double> d_ms32;               // FP32 mode: mean and scale of every element, [2][n_elem][NG]
  bool g32_dirty = true;
  // normalised FP64 copy of the descriptors, (x - mean) * (1 / scale) per atom's element: the training kernels then skip
  // the per-tile normalisation pass (NN.cpp:647-651).  Refreshed when descriptors or input scalings change.
  DevBuf<double> d_Gn, d_msn;
  bool gn_dirty = true, gn_ready = false;
  bool fp32_umma = false;              // FP32 epochs run the tcgen05 kernel (decided with the tile list)
  bool fp32_umma_ran = false;          // a tcgen05 FP32 launch is in flight / unchecked (its error flag is read at the next host sync)
  cudaStream_t stream = nullptr;
  long long launches = 0;
  double phase_ms[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  bool epoch_phases_pending = false;   // pp_epoch finished: events 0..3 hold its phases, read out on demand (pp_last_phase_ms)
  cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};

  bool spec_set = false;
  DescSpec spec;
  DescSpec* d_spec = nullptr;
  DevBuf<int> d_overflow;

  NetHost nets[PP_MAX_ELEM];
  int P_total = 0;
  DevBuf<double> d_params;          // all elements, concatenated
  DevBuf<double> d_paramsT;         // the same with every weight layer transposed ([n_in + 1][n_out]): fused MD kernel
  bool paramsT_dirty = true;
  double out_mean = 0.0, out_scale = 1.0;
  DevBuf<double> d_tanh_table;

  // ---- training set ----
  int n_struct = 0;
  long long n_atoms = 0;
  int desc_cap = 64;                   // neighbour capacity per atom: doubled (and the launch repeated) when an atom overflows it
  std::vector<long long> h_struct_atom0{0};
  std::vector<double> h_FtoA, h_target;
  std::vector<int> h_mxyz, h_tag, h_elem, h_atom_struct;
  std::vector<double> h_frac;
  std::vector<int> h_rows[PP_MAX_ELEM];
  DevBuf<long long> d_struct_atom0;
  DevBuf<double> d_FtoA, d_target, d_frac, d_G, d_E_atom, d_E_struct, d_cstruct;
  DevBuf<int> d_mxyz, d_tag, d_elem, d_atom_struct;
  DevBuf<int> d_rows[PP_MAX_ELEM];
  DevBuf<double> d_partial, d_result, d_stat_partial, d_stat_sums;
  int epoch_form = -1;                 // kernel form of the last epoch (pp_epoch_form)
  DevBuf<double> d_err_scratch;        // struct_errors_kernel: per-block loss terms
  DevBuf<double> d_scratch2;           // mlp_fused2_kernel: derivative scratch (stays in L2)
  DevBuf<double> d_hstore[PP_MAX_ELEM], d_tstore[PP_MAX_ELEM];   // stored two-phase epoch: H_l and T'_l of every atom of an element
  DevBuf<unsigned> d_err_ticket;
  DevBuf<long long> d_count;
  double* h_result = nullptr;          // pinned mirror of d_result ([P_total + 3]) and of the parameters ([P_total]) behind it
  size_t h_result_cap = 0;
  // fused single-pass epoch: tiles of whole structures
  bool tiles_dirty = true, fused_ok = false;
  int tiles_tm = 0;                    // tile size the list was packed for
  int n_fused_tiles = 0;
  DevBuf<int> d_tile_struct0;

  // ---- MD ----
  std::vector<int> type_to_elem;
  int md_cap = 64;
  int md_nlocal = 0, md_ntotal = 0;
  long long md_nneigh = 0;
  DevBuf<double> d_x, d_f, d_Gmd, d_dEdG, d_eatom;
  DevBuf<int> d_md_elem, d_neigh;
  DevBuf<long long> d_neigh_off;
  DevBuf<int> d_md_rows[PP_MAX_ELEM];
  int md_rows_n[PP_MAX_ELEM] = {0, 0, 0, 0};
  std::vector<int> h_md_elem;
  // cell-list work space and resident-step buffers
  DevBuf<int> d_cell_of, d_cell_count, d_sorted, d_counts;
  DevBuf<long long> d_cell_start, d_scan_sums, d_scan_total;
  DevBuf<double> d_sum_partial, d_etotal;
  bool md_list_resident = false;
  bool md_fused_last = false;          // the last MD step ran the single fused kernel (its time is phase 6)
  // reverse communication of ghost forces (pp_md_exchange_*)
  struct SegList { int n_seg = 0; DevBuf<int> seg_off, seg_dst, idx; };
  struct MdPeer { int rank = 0; int n_send = 0, n_recv = 0; DevBuf<int> send_idx; DevBuf<double> sendbuf, recvbuf; SegList add; };
  SegList md_self;
  std::vector<MdPeer> md_peers;
  bool md_exchange_set = false;

  PPComm comm;
};

static int tick(pp_context* c, int i)
{
  if (i == 0) c->epoch_phases_pending = false;      // a new timed operation re-records the events
  PP_CUDA(cudaEventRecord(c->ev[i], c->stream));
  return 0;
}
static int tock(pp_context* c, int i0, int i1, int which)
{
  PP_CUDA(cudaEventSynchronize(c->ev[i1]));
  float ms = 0;
  PP_CUDA(cudaEventElapsedTime(&ms, c->ev[i0], c->ev[i1]));
  c->phase_ms[which] = ms;
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
extern "C" int pp_create(int device, int fp_mode, pp_context** out)
{
  PP_REQUIRE(out != nullptr, "pp_create: out is NULL");
  PP_REQUIRE(fp_mode == 64 || fp_mode == 32, "pp_create: fp_mode %d not implemented (64 or 32)", fp_mode);
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  PP_REQUIRE(e == cudaSuccess && ndev > 0, "pp_create: no CUDA device available (%s); this library has no CPU fallback",
             e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
  PP_REQUIRE(device >= 0 && device < ndev, "pp_create: device %d out of range (have %d)", device, ndev);
  PP_CUDA(cudaSetDevice(device));
  pp_context* c = new pp_context();
  c->device = device;
  c->fp_mode = fp_mode;
  PP_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  for (int i = 0; i < 6; i++) PP_CUDA(cudaEventCreate(&c->ev[i]));
  PP_CUDA(cudaMalloc(&c->d_spec, sizeof(DescSpec)));
  if (c->d_overflow.ensure(1, 0, c->stream)) return 1;
  std::vector<double> table;
  ppk_make_tanh_table(table);
  if (c->d_tanh_table.ensure(table.size(), 0, c->stream)) return 1;
  PP_CUDA(cudaMemcpy(c->d_tanh_table.p, table.data(), table.size() * sizeof(double), cudaMemcpyHostToDevice));
  *out = c;
  return 0;
}

extern "C" void pp_destroy(pp_context* c)
{
  if (!c) return;
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  c->comm.destroy();
  // DevBuf members are plain pointers: release explicitly
  c->d_overflow.release(); c->d_params.release(); c->d_paramsT.release(); c->d_tanh_table.release();
  c->d_struct_atom0.release(); c->d_FtoA.release(); c->d_target.release(); c->d_frac.release(); c->d_G.release();
  c->d_E_atom.release(); c->d_E_struct.release(); c->d_cstruct.release(); c->d_mxyz.release(); c->d_tag.release();
  c->d_elem.release(); c->d_Gn.release(); c->d_msn.release(); c->d_atom_struct.release(); c->d_partial.release(); c->d_result.release(); c->d_err_scratch.release(); c->d_scratch2.release(); c->d_err_ticket.release();
  if (c->h_result) { cudaFreeHost(c->h_result); c->h_result = nullptr; c->h_result_cap = 0; }

  c->d_stat_partial.release(); c->d_stat_sums.release(); c->d_count.release();
  c->d_x.release(); c->d_f.release(); c->d_Gmd.release(); c->d_dEdG.release(); c->d_eatom.release();
  c->d_md_elem.release(); c->d_neigh.release(); c->d_neigh_off.release();
  c->d_cell_of.release(); c->d_cell_count.release(); c->d_sorted.release(); c->d_counts.release();
  c->d_cell_start.release(); c->d_scan_sums.release(); c->d_scan_total.release(); c->d_sum_partial.release(); c->d_etotal.release();
  for (int e = 0; e < PP_MAX_ELEM; e++) {
    c->d_rows[e].release(); c->d_md_rows[e].release(); c->d_hstore[e].release(); c->d_tstore[e].release();
    c->nets[e].d_mean.release(); c->nets[e].d_scale.release(); c->nets[e].d_tf.release(); c->nets[e].d_dw_tiles.release(); c->nets[e].d_dw_tiles2.release(); c->nets[e].d_dw_tiles3.release(); c->nets[e].d_dw_tiles32.release();
  }
  c->d_tile_struct0.release(); c->d_G32.release(); c->d_ms32.release();
  c->md_self.seg_off.release(); c->md_self.seg_dst.release(); c->md_self.idx.release();
  for (auto& p : c->md_peers) {
    p.send_idx.release(); p.sendbuf.release(); p.recvbuf.release();
    p.add.seg_off.release(); p.add.seg_dst.release(); p.add.idx.release();
  }
  if (c->d_spec) cudaFree(c->d_spec);
  for (int i = 0; i < 6; i++) if (c->ev[i]) cudaEventDestroy(c->ev[i]);
  if (c->stream) cudaStreamDestroy(c->stream);
  delete c;
}

extern "C" long long pp_kernel_launches(pp_context* c) { return c ? c->launches : 0; }
extern "C" double pp_last_phase_ms(pp_context* c, int which)
{
  if (!c || which < 0 || which >= 8) return 0.0;
  if (which == 7 && c->md_exchange_set) tock(c, 4, 5, 7);      // the exchange's own pair of events (synchronises on them)
  if (c->epoch_phases_pending) {       // the epoch's events are only turned into times when somebody asks
    c->epoch_phases_pending = false;
    tock(c, 0, 1, 1); tock(c, 1, 2, 2); tock(c, 2, 3, 3);
  }
  return c->phase_ms[which];
}
extern "C" void* pp_stream(pp_context* c) { return c ? (void*)c->stream : nullptr; }

// ---------------------------------------------------------------------------------------------------------
// Structure::init_G / Allocate_G / init_type_G_map (Structure.cpp:779-960, 1491-1512)
extern "C" int pp_set_descriptors(pp_context* c, int n_elem, int nG1, const double* G1, int nG2, const double* G2,
                                  int nG3, const double* G3, int nG4, const double* G4, double Rcut)
{
  PP_REQUIRE(c, "null context");
  PP_REQUIRE(n_elem >= 1 && n_elem <= PP_MAX_ELEM, "pp_set_descriptors: %d elements (max %d)", n_elem, PP_MAX_ELEM);
  PP_REQUIRE(nG1 >= 0 && nG2 >= 0 && nG3 >= 0 && nG4 >= 0, "pp_set_descriptors: negative row count");
  PP_REQUIRE(nG1 + nG2 <= PP_MAX_PAIR_ROWS, "pp_set_descriptors: %d pair rows (max %d)", nG1 + nG2, PP_MAX_PAIR_ROWS);
  PP_REQUIRE(nG4 <= PP_MAX_G4_ROWS, "pp_set_descriptors: %d G4 rows (max %d)", nG4, PP_MAX_G4_ROWS);
  PP_REQUIRE(nG3 <= PP_MAX_G3_ROWS, "pp_set_descriptors: %d G3 rows (max %d)", nG3, PP_MAX_G3_ROWS);
  PP_REQUIRE(Rcut > 0, "pp_set_descriptors: Rcut must be positive");
  // with structures resident only the row PARAMETERS may change (the -optimize search, Potential.cpp:709-877):
  // same element count, row counts and Rcut, so every buffer keeps its shape; pp_recompute_descriptors then refreshes
  if (c->n_struct > 0)
    PP_REQUIRE(c->spec_set && c->spec.n_elem == n_elem && c->spec.nG1 == nG1 && c->spec.nG2 == nG2 && c->spec.nG3 == nG3 &&
                   c->spec.nG4 == nG4 && c->spec.Rcut == Rcut,
               "pp_set_descriptors: structures already added; only row parameters may change (same counts and Rcut), "
               "otherwise call pp_clear_structures first");
  PP_CUDA(cudaSetDevice(c->device));
  // the new specification is assembled in a local and committed (host copy, device copy, flags) only after every check
  // passed: a rejected call leaves the context as it was
  DescSpec S_new;
  DescSpec& S = S_new;
  memset(&S, 0, sizeof(S));
  S.n_elem = n_elem; S.nG1 = nG1; S.nG2 = nG2; S.nG3 = nG3; S.nG4 = nG4;
  S.npair = nG1 + nG2; S.ntrip = nG3 + nG4;
  S.NG = n_elem * S.npair + n_elem * (n_elem + 1) / 2 * S.ntrip;
  S.Rcut = Rcut;
  for (int r = 0; r < nG1; r++) {
    S.pair_rc[r] = G1[2 * r]; S.pair_fc[r] = (int)G1[2 * r + 1]; S.pair_eta[r] = 0; S.pair_rs[r] = 0;
  }
  for (int r = 0; r < nG2; r++) {
    S.pair_rc[nG1 + r] = G2[4 * r]; S.pair_fc[nG1 + r] = (int)G2[4 * r + 1];
    S.pair_eta[nG1 + r] = G2[4 * r + 2]; S.pair_rs[nG1 + r] = G2[4 * r + 3];
  }
  for (int r = 0; r < nG3; r++) {
    S.g3_rc[r] = G3[5 * r]; S.g3_fc[r] = (int)G3[5 * r + 1]; S.g3_eta[r] = G3[5 * r + 2];
    S.g3_xi[r] = G3[5 * r + 3]; S.g3_lam[r] = G3[5 * r + 4];
  }
  // factorise the G4 rows into radial keys (Rc, fc, eta) and angular keys (xi, lambda), first-seen order
  S.NR = 0; S.NA = 0;
  for (int r = 0; r < nG4; r++) {
    const double* row = G4 + 5 * r;
    int ri = -1, ai = -1;
    for (int k = 0; k < S.NR; k++)
      if (S.rad_rc[k] == row[0] && S.rad_fc[k] == (int)row[1] && S.rad_eta[k] == row[2]) ri = k;
    if (ri < 0) {
      PP_REQUIRE(S.NR < PP_MAX_RAD, "pp_set_descriptors: more than %d distinct (Rc, fc, eta) among G4 rows", PP_MAX_RAD);
      ri = S.NR++;
      S.rad_rc[ri] = row[0]; S.rad_fc[ri] = (int)row[1]; S.rad_eta[ri] = row[2];
    }
    for (int k = 0; k < S.NA; k++)
      if (S.ang_xi[k] == row[3] && S.ang_lam[k] == row[4]) ai = k;
    if (ai < 0) {
      PP_REQUIRE(S.NA < PP_MAX_ANG, "pp_set_descriptors: more than %d distinct (xi, lambda) among G4 rows", PP_MAX_ANG);
      ai = S.NA++;
      S.ang_xi[ai] = row[3]; S.ang_lam[ai] = row[4];
      S.ang_pref[ai] = pow(2.0, 1.0 - row[3]);                         // Structure.cpp:811-813
      bool is_int = (row[3] == (double)(int)row[3]) && row[3] >= 0;     // Structure.h:257 (negative: 1/base path)
      S.ang_xi_int[ai] = is_int ? (int)row[3] : -1;
      PP_REQUIRE(row[3] >= 0 || !(row[3] == (double)(int)row[3]), "pp_set_descriptors: negative integer xi not supported");
    }
    S.g4_rad[r] = ri; S.g4_ang[r] = ai;
  }
  if (S.NR == 0) S.NR = 1;   // keeps shared-memory strides non-degenerate when there are no G4 rows
  // Order the angular keys so that powers of two of the same lambda form chains x -> x^2 -> x^4 ... and build the
  // evaluation program (pp_desc.cuh, AngKeys): a chain key costs only its squarings relative to the previous key.
  {
    const int NA = S.NA;
    std::vector<int> order(NA);
    for (int k = 0; k < NA; k++) order[k] = k;
    auto is_pow2 = [&](int k) { int e = S.ang_xi_int[k]; return e >= 1 && (e & (e - 1)) == 0 && e <= 1024; };
    std::stable_sort(order.begin(), order.end(), [&](int x, int y) {
      bool px = is_pow2(x), py = is_pow2(y);
      if (px != py) return px;
      if (S.ang_lam[x] != S.ang_lam[y]) return S.ang_lam[x] > S.ang_lam[y];
      return S.ang_xi[x] < S.ang_xi[y];
    });
    DescSpec T = S;
    std::vector<int> newpos(NA);
    for (int k = 0; k < NA; k++) {
      int o = order[k];
      newpos[o] = k;
      S.ang_xi[k] = T.ang_xi[o]; S.ang_lam[k] = T.ang_lam[o]; S.ang_pref[k] = T.ang_pref[o]; S.ang_xi_int[k] = T.ang_xi_int[o];
    }
    for (int r = 0; r < nG4; r++) S.g4_ang[r] = newpos[T.g4_ang[r]];
    auto ilog2 = [](int v) { int l = 0; while (v > 1) { v >>= 1; l++; } return l; };
    for (int k = 0; k < NA; k++) {
      int e = S.ang_xi_int[k];
      int prog;
      if (e < 0) prog = 3;
      else if (!(e >= 1 && (e & (e - 1)) == 0 && e <= 1024)) prog = 2;
      else {
        bool fresh = (k % 4 == 0) || (S.ang_prog[k - 1] & 3) != 1 || S.ang_lam[k - 1] != S.ang_lam[k];
        int ns = fresh ? ilog2(e) : ilog2(e / S.ang_xi_int[k - 1]);
        if (ns > 5) prog = 2;
        else prog = 1 | (fresh ? 4 : 0) | (ns << 3);
      }
      S.ang_prog[k] = prog;
    }
  }
  // power-basis fast path: every xi an integer power of two below 2^5 and at most two distinct lambdas
  {
    S.basis_nb = 0;
    S.basis_used[0] = S.basis_used[1] = 0;
    S.basis_lam[0] = S.basis_lam[1] = 0.0;
    bool ok = nG4 > 0;
    int nlam = 0, maxb = 0;
    for (int k = 0; k < S.NA && ok; k++) {
      int e = S.ang_xi_int[k];
      if (!(e >= 1 && (e & (e - 1)) == 0 && e <= 16)) { ok = false; break; }
      int b = 0;
      while ((1 << b) < e) b++;
      maxb = std::max(maxb, b);
      int g = -1;
      for (int q = 0; q < nlam; q++) if (S.basis_lam[q] == S.ang_lam[k]) g = q;
      if (g < 0) {
        if (nlam == 2) { ok = false; break; }
        S.basis_lam[nlam++] = S.ang_lam[k];
      }
    }
    if (ok) {
      S.basis_nb = maxb + 1 <= 3 ? 3 : 5;
      for (int r = 0; r < nG4; r++) {
        int k = S.g4_ang[r];
        int e = S.ang_xi_int[k], b = 0;
        while ((1 << b) < e) b++;
        S.g4_bb[r] = b;
        S.g4_bg[r] = (S.basis_lam[0] == S.ang_lam[k]) ? 0 : 1;
        S.basis_used[S.g4_bg[r]] |= 1 << b;
      }
    }
  }
  // evaluation order of the Gaussian factors (DescSpec::pair_order / rad_order): group by Rs, ascending eta, and chain an
  // eta that is an integer multiple (<= 64) of its predecessor's; at most a 4096-th power of a computed exponential in all
  {
    auto chain_of = [](double eta, double eta_prev, long long& cum) {
      if (!(eta_prev > 0.0) || !(eta > 0.0)) return 0;
      const double m = nearbyint(eta / eta_prev);
      if (m < 1.0 || m > 64.0 || fabs(eta - m * eta_prev) > 4e-16 * eta) return 0;
      if (cum * (long long)m > 4096) return 0;
      cum *= (long long)m;
      return (int)m;
    };
    std::vector<int> ord;
    for (int r = nG1; r < S.npair; r++) ord.push_back(r);
    std::stable_sort(ord.begin(), ord.end(), [&](int x, int y) {
      if (S.pair_rs[x] != S.pair_rs[y]) return S.pair_rs[x] < S.pair_rs[y];
      return S.pair_eta[x] < S.pair_eta[y];
    });
    for (int r = 0; r < nG1; r++) { S.pair_order[r] = r; S.pair_chain[r] = 0; }
    long long cum = 1;
    for (size_t k = 0; k < ord.size(); k++) {
      const int r = ord[k];
      int m = 0;
      if (k > 0 && S.pair_rs[ord[k - 1]] == S.pair_rs[r]) m = chain_of(S.pair_eta[r], S.pair_eta[ord[k - 1]], cum);
      if (m == 0) cum = 1;
      S.pair_order[nG1 + k] = r; S.pair_chain[nG1 + k] = m;
    }
    ord.clear();
    for (int r = 0; r < S.NR; r++) ord.push_back(r);
    std::stable_sort(ord.begin(), ord.end(), [&](int x, int y) { return S.rad_eta[x] < S.rad_eta[y]; });
    cum = 1;
    for (size_t k = 0; k < ord.size(); k++) {
      int m = 0;
      if (k > 0) m = chain_of(S.rad_eta[ord[k]], S.rad_eta[ord[k - 1]], cum);
      if (m == 0) cum = 1;
      S.rad_order[k] = ord[k]; S.rad_chain[k] = m;
    }
  }
  // squared-distance thresholds equivalent to the reference's tests on d = sqrt(d2) (Structure.cpp:1337)
  {
    auto thr = [](double r) {
      double y = r * r;
      while (sqrt(y) > r) y = nextafter(y, 0.0);
      while (sqrt(nextafter(y, INFINITY)) <= r) y = nextafter(y, INFINITY);
      return y;
    };
    S.scan_d2_hi = thr(Rcut);
    S.scan_d2_lo = thr(0.01);
  }
  int index = 0;
  for (int i = 0; i < n_elem; i++)
    for (int j = i; j < n_elem; j++) {
      S.trip_base[i][j] = S.trip_base[j][i] = S.npair * n_elem + index;
      index += S.ntrip;
    }
  PP_CUDA(cudaMemcpy(c->d_spec, &S, sizeof(S), cudaMemcpyHostToDevice));
  if (c->spec_set && c->spec.NG != S.NG)
    for (int e = 0; e < PP_MAX_ELEM; e++) c->nets[e].set = false;      // networks of the old input width no longer apply
  c->spec = S;
  c->spec_set = true;
  c->g32_dirty = true;
  c->gn_dirty = true;
  return 0;
}

extern "C" int pp_num_descriptors(pp_context* c) { return (c && c->spec_set) ? c->spec.NG : -1; }

// ---------------------------------------------------------------------------------------------------------
static int rebuild_param_layout(pp_context* c)
{
  int off = 0;
  for (int e = 0; e < PP_MAX_ELEM; e++) {
    if (!c->nets[e].set) continue;
    c->nets[e].p_off = off;
    off += c->nets[e].P;
  }
  c->P_total = off;
  c->paramsT_dirty = true;
  if (c->d_params.ensure((size_t)std::max(off, 1), 0, c->stream)) return 1;
  for (int e = 0; e < PP_MAX_ELEM; e++) {
    NetHost& n = c->nets[e];
    if (!n.set) continue;
    PP_CUDA(cudaMemcpyAsync(c->d_params.p + n.p_off, n.params.data(), n.P * sizeof(double), cudaMemcpyHostToDevice, c->stream));
    n.dev.params = c->d_params.p + n.p_off;
  }
  PP_CUDA(cudaStreamSynchronize(c->stream));
  return 0;
}

extern "C" int pp_set_network(pp_context* c, int elem, int n_layers, const int* sizes, const int* node_transfer,
                              const double* params, const double* in_mean, const double* in_scale)
{
  PP_REQUIRE(c && c->spec_set, "pp_set_network: call pp_set_descriptors first");
  PP_CUDA(cudaSetDevice(c->device));
  PP_REQUIRE(elem >= 0 && elem < c->spec.n_elem, "pp_set_network: element index %d out of range", elem);
  PP_REQUIRE(n_layers >= 1 && n_layers <= PP_MAX_LAYERS, "pp_set_network: %d weight layers (max %d)", n_layers, PP_MAX_LAYERS);
  PP_REQUIRE(sizes[0] == c->spec.NG, "pp_set_network: input width %d != number of descriptors %d", sizes[0], c->spec.NG);
  PP_REQUIRE(sizes[n_layers] == 1, "pp_set_network: the output layer must have one node");
  NetHost& n = c->nets[elem];
  n.n_layers = n_layers;
  n.sizes.assign(sizes, sizes + n_layers + 1);
  n.P = 0; n.n_nodes = 0;
  memset(&n.dev, 0, sizeof(n.dev));
  n.dev.n_layers = n_layers;
  for (int l = 0; l <= n_layers; l++) n.dev.sizes[l] = sizes[l];
  for (int l = 0; l < n_layers; l++) {
    PP_REQUIRE(sizes[l + 1] >= 1, "pp_set_network: empty layer");
    n.dev.w_off[l] = n.P;
    n.dev.node_off[l] = n.n_nodes;
    n.P += sizes[l + 1] * (sizes[l] + 1);
    n.n_nodes += sizes[l + 1];
  }
  n.dev.P = n.P;
  n.tf.assign(node_transfer, node_transfer + n.n_nodes);
  for (int i = 0; i < n.n_nodes; i++) PP_REQUIRE(n.tf[i] >= 0 && n.tf[i] <= 8, "pp_set_network: unknown transfer id %d", n.tf[i]);
  PP_REQUIRE(n.tf[n.n_nodes - 1] == PP_TF_LINEAR, "pp_set_network: the output node must be linear (NN.cpp:76)");
  n.params.assign(params, params + n.P);
  n.mean.assign(in_mean, in_mean + sizes[0]);
  n.scale.assign(in_scale, in_scale + sizes[0]);
  if (n.d_mean.ensure(sizes[0], 0, c->stream) || n.d_scale.ensure(sizes[0], 0, c->stream) || n.d_tf.ensure(n.n_nodes, 0, c->stream)) return 1;
  PP_CUDA(cudaMemcpy(n.d_mean.p, n.mean.data(), sizes[0] * sizeof(double), cudaMemcpyHostToDevice));
  PP_CUDA(cudaMemcpy(n.d_scale.p, n.scale.data(), sizes[0] * sizeof(double), cudaMemcpyHostToDevice));
  PP_CUDA(cudaMemcpy(n.d_tf.p, n.tf.data(), n.n_nodes * sizeof(int), cudaMemcpyHostToDevice));
  n.dev.mean = n.d_mean.p; n.dev.scale = n.d_scale.p; n.dev.tf = n.d_tf.p;
  n.dev.tanh_layers = 0;
  n.dev.spline_layers = 0;
  for (int l = 0; l < n_layers - 1; l++) {
    bool all = true;
    for (int k = 0; k < sizes[l + 1]; k++) all = all && n.tf[n.dev.node_off[l] + k] == 1;
    if (all) n.dev.tanh_layers |= 1u << (l + 1);
    bool all_spline = true;
    for (int k = 0; k < sizes[l + 1]; k++) all_spline = all_spline && n.tf[n.dev.node_off[l] + k] == 2;
    if (all_spline) n.dev.spline_layers |= 1u << (l + 1);
  }
  {
    n.tile_atoms = ppk_mlp_tile_atoms(n.dev);
    PP_REQUIRE(n.tile_atoms > 0, "pp_set_network: network too large for shared memory");
    int nt = ppk_mlp_num_dw_tiles(n.dev, nullptr);
    std::vector<unsigned char> tiles((size_t)std::max(nt, 1) * ppk_mlp_dw_tile_bytes());
    ppk_mlp_num_dw_tiles(n.dev, tiles.data());
    if (n.d_dw_tiles.ensure(tiles.size(), 0, c->stream)) return 1;
    PP_CUDA(cudaMemcpy(n.d_dw_tiles.p, tiles.data(), tiles.size(), cudaMemcpyHostToDevice));
    // work table of the two-CTA fused training kernel, when this network is eligible for it
    int n2 = 0;
    n.has_tiles2 = false;
    if (ppk_mlp2_table(n.dev, nullptr, &n2) && n2 > 0) {
      std::vector<unsigned char> t2((size_t)n2 * ppk_mlp_dw_tile_bytes());
      ppk_mlp2_table(n.dev, t2.data(), &n2);
      if (n.d_dw_tiles2.ensure(t2.size(), 0, c->stream)) return 1;
      PP_CUDA(cudaMemcpy(n.d_dw_tiles2.p, t2.data(), t2.size(), cudaMemcpyHostToDevice));
      n.has_tiles2 = true;
    }
    int n32 = 0;
    n.has_tiles32 = false;
    const int esz = ppk_mlp32_table(n.dev, nullptr, &n32);
    if (esz > 0 && n32 > 0) {
      std::vector<unsigned char> t32((size_t)n32 * esz);
      ppk_mlp32_table(n.dev, t32.data(), &n32);
      if (n.d_dw_tiles32.ensure(t32.size(), 0, c->stream)) return 1;
      PP_CUDA(cudaMemcpy(n.d_dw_tiles32.p, t32.data(), t32.size(), cudaMemcpyHostToDevice));
      n.has_tiles32 = true;
    }
    int n3 = 0;
    n.has_tiles3 = false;
    if (ppk_mlpw_table(n.dev, nullptr, &n3) && n3 > 0) {
      std::vector<unsigned char> t3((size_t)n3 * ppk_mlp_dw_tile_bytes());
      ppk_mlpw_table(n.dev, t3.data(), &n3);
      if (n.d_dw_tiles3.ensure(t3.size(), 0, c->stream)) return 1;
      PP_CUDA(cudaMemcpy(n.d_dw_tiles3.p, t3.data(), t3.size(), cudaMemcpyHostToDevice));
      n.has_tiles3 = true;
    }
  }
  c->tiles_dirty = true;
  c->g32_dirty = true;
  c->gn_dirty = true;
  n.set = true;
  return rebuild_param_layout(c);
}

extern "C" int pp_set_output_scaling(pp_context* c, double out_mean, double out_scale)
{
  PP_REQUIRE(c, "null context");
  c->out_mean = out_mean;
  c->out_scale = out_scale;
  return 0;
}

extern "C" int pp_num_parameters(pp_context* c) { return c ? c->P_total : -1; }

static int nets_ready(pp_context* c)
{
  PP_REQUIRE(c && c->spec_set, "descriptors not set");
  for (int e = 0; e < c->spec.n_elem; e++) PP_REQUIRE(c->nets[e].set, "no network set for element index %d", e);
  return 0;
}

extern "C" int pp_set_parameters(pp_context* c, const double* params)
{
  if (nets_ready(c)) return 1;
  PP_CUDA(cudaSetDevice(c->device));
  c->paramsT_dirty = true;
  for (int e = 0; e < c->spec.n_elem; e++) {
    NetHost& n = c->nets[e];
    std::copy(params + n.p_off, params + n.p_off + n.P, n.params.begin());
  }
  PP_CUDA(cudaMemcpyAsync(c->d_params.p, params, c->P_total * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  PP_CUDA(cudaStreamSynchronize(c->stream));
  return 0;
}

extern "C" int pp_get_parameters(pp_context* c, double* params)
{
  if (nets_ready(c)) return 1;
  PP_CUDA(cudaSetDevice(c->device));
  PP_CUDA(cudaMemcpyAsync(params, c->d_params.p, c->P_total * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  PP_CUDA(cudaStreamSynchronize(c->stream));
  return 0;
}

extern "C" int pp_set_input_scaling(pp_context* c, int elem, const double* in_mean, const double* in_scale)
{
  PP_REQUIRE(c && elem >= 0 && elem < PP_MAX_ELEM && c->nets[elem].set, "pp_set_input_scaling: no network for element %d", elem);
  PP_CUDA(cudaSetDevice(c->device));
  NetHost& n = c->nets[elem];
  int NG = n.sizes[0];
  n.mean.assign(in_mean, in_mean + NG);
  n.scale.assign(in_scale, in_scale + NG);
  c->g32_dirty = true;
  c->gn_dirty = true;
  PP_CUDA(cudaMemcpy(n.d_mean.p, n.mean.data(), NG * sizeof(double), cudaMemcpyHostToDevice));
  PP_CUDA(cudaMemcpy(n.d_scale.p, n.scale.data(), NG * sizeof(double), cudaMemcpyHostToDevice));
  return 0;
}

extern "C" int pp_get_input_scaling(pp_context* c, int elem, double* in_mean, double* in_scale)
{
  PP_REQUIRE(c && elem >= 0 && elem < PP_MAX_ELEM && c->nets[elem].set, "pp_get_input_scaling: no network for element %d", elem);
  NetHost& n = c->nets[elem];
  std::copy(n.mean.begin(), n.mean.end(), in_mean);
  std::copy(n.scale.begin(), n.scale.end(), in_scale);
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
// Structure::Set_Convert + Normalize (Structure.cpp:1159-1279), host arithmetic identical to the reference
// (this translation unit is compiled with -ffp-contract=off for host code).
extern "C" int pp_cart_to_frac(const double* cell, const double* cart, int n, double* frac)
{
  PP_REQUIRE(cell && cart && frac && n >= 0, "pp_cart_to_frac: bad arguments");
  // FtoA = transpose(cell); AtoF = inverse by cofactors (Structure.h:180-197)
  double m[3][3];
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) m[j][i] = cell[3 * i + j];
  double det = m[0][0] * (m[1][1] * m[2][2] - m[2][1] * m[1][2]) - m[0][1] * (m[1][0] * m[2][2] - m[1][2] * m[2][0]) +
                        m[0][2] * (m[1][0] * m[2][1] - m[1][1] * m[2][0]);
  PP_REQUIRE(det != 0, "Non-invertible unit cell");
  double inv = 1 / det;
  double A[3][3];
  A[0][0] = (m[1][1] * m[2][2] - m[2][1] * m[1][2]) * inv;
  A[0][1] = (m[0][2] * m[2][1] - m[0][1] * m[2][2]) * inv;
  A[0][2] = (m[0][1] * m[1][2] - m[0][2] * m[1][1]) * inv;
  A[1][0] = (m[1][2] * m[2][0] - m[1][0] * m[2][2]) * inv;
  A[1][1] = (m[0][0] * m[2][2] - m[0][2] * m[2][0]) * inv;
  A[1][2] = (m[1][0] * m[0][2] - m[0][0] * m[1][2]) * inv;
  A[2][0] = (m[1][0] * m[2][1] - m[2][0] * m[1][1]) * inv;
  A[2][1] = (m[2][0] * m[0][1] - m[0][0] * m[2][1]) * inv;
  A[2][2] = (m[0][0] * m[1][1] - m[1][0] * m[0][1]) * inv;
  for (int i = 0; i < n; i++) {
    const double* p = cart + 3 * (size_t)i;
    for (int r = 0; r < 3; r++) {
      double v = A[r][0] * p[0] + A[r][1] * p[1] + A[r][2] * p[2];
      if (v < 0) v += 1;
      else if (v > 1) v -= 1;
      frac[3 * (size_t)i + r] = v;
    }
  }
  return 0;
}

// The same for a batch of structures in one call (the Potential constructor's loop over systems, Potential.cpp:113-150):
// structure s has natoms[s] atoms, cell cells[9 s ..], positions concatenated.
extern "C" int pp_cart_to_frac_batch(int n_struct, const int* natoms, const double* cells, const double* cart, double* frac)
{
  PP_REQUIRE(n_struct >= 0 && (n_struct == 0 || (natoms && cells && cart && frac)), "pp_cart_to_frac_batch: bad arguments");
  size_t a0 = 0;
  for (int s = 0; s < n_struct; s++) {
    PP_REQUIRE(natoms[s] >= 0, "pp_cart_to_frac_batch: structure %d has %d atoms", s, natoms[s]);
    if (pp_cart_to_frac(cells + 9 * (size_t)s, cart + 3 * a0, natoms[s], frac + 3 * a0)) return 1;
    a0 += (size_t)natoms[s];
  }
  return 0;
}

static PeriodicArgs periodic_args(pp_context* c)
{
  PeriodicArgs a;
  memset(&a, 0, sizeof(a));
  a.spec = c->d_spec;
  a.atom_struct = c->d_atom_struct.p;
  a.struct_atom0 = c->d_struct_atom0.p;
  a.FtoA = c->d_FtoA.p;
  a.mxyz = c->d_mxyz.p;
  a.frac = c->d_frac.p;
  a.elem = c->d_elem.p;
  a.G = c->d_G.p;
  a.overflow = c->d_overflow.p;
  return a;
}

extern "C" int pp_add_structures(pp_context* c, int n_struct, const int* natoms, const double* cells, const double* frac,
                                 const int* elem, const double* targets, const int* tags)
{
  PP_REQUIRE(c && c->spec_set, "pp_add_structures: call pp_set_descriptors first");
  PP_REQUIRE(n_struct >= 0, "pp_add_structures: negative count");
  if (n_struct == 0) return 0;
  PP_CUDA(cudaSetDevice(c->device));
  const DescSpec& S = c->spec;
  const int s0 = c->n_struct;
  const long long a0 = c->n_atoms;
  long long add_atoms = 0;
  for (int s = 0; s < n_struct; s++) {
    PP_REQUIRE(natoms[s] > 0, "pp_add_structures: structure %d has %d atoms", s, natoms[s]);
    add_atoms += natoms[s];
  }
  for (long long i = 0; i < add_atoms; i++)
    PP_REQUIRE(elem[i] >= 0 && elem[i] < S.n_elem, "pp_add_structures: element index %d out of range", elem[i]);
  // host mirrors
  c->h_frac.insert(c->h_frac.end(), frac, frac + 3 * add_atoms);
  c->h_elem.insert(c->h_elem.end(), elem, elem + add_atoms);
  for (int s = 0; s < n_struct; s++) {
    const double* cell = cells + 9 * (size_t)s;
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) c->h_FtoA.push_back(cell[3 * j + i]);   // FtoA = cell^T
    // Structure.cpp:1300-1302 with the norms of Set_Convert (:1168-1173)
    for (int v = 0; v < 3; v++) {
      double nrm = sqrt(pow(cell[3 * v], 2) + pow(cell[3 * v + 1], 2) + pow(cell[3 * v + 2], 2));
      PP_REQUIRE(nrm > 0, "pp_add_structures: zero lattice vector");
      c->h_mxyz.push_back((int)ceil(1.1 * S.Rcut / nrm));
    }
    c->h_target.push_back(targets[s]);
    c->h_tag.push_back(tags[s]);
    for (int i = 0; i < natoms[s]; i++) c->h_atom_struct.push_back(s0 + s);
    c->h_struct_atom0.push_back(c->h_struct_atom0.back() + natoms[s]);
  }
  const int ns = s0 + n_struct;
  const long long na = a0 + add_atoms;
  cudaStream_t st = c->stream;
  if (c->d_struct_atom0.ensure(ns + 1, s0 + 1, st) || c->d_FtoA.ensure(9 * (size_t)ns, 9 * (size_t)s0, st) ||
      c->d_mxyz.ensure(3 * (size_t)ns, 3 * (size_t)s0, st) || c->d_target.ensure(ns, s0, st) || c->d_tag.ensure(ns, s0, st) ||
      c->d_frac.ensure(3 * (size_t)na, 3 * (size_t)a0, st) || c->d_elem.ensure(na, a0, st) ||
      c->d_atom_struct.ensure(na, a0, st) || c->d_G.ensure((size_t)na * S.NG, (size_t)a0 * S.NG, st) ||
      c->d_E_atom.ensure(na, 0, st) || c->d_E_struct.ensure(ns, 0, st) || c->d_cstruct.ensure(ns, 0, st))
    return 1;
  PP_CUDA(cudaMemcpyAsync(c->d_struct_atom0.p, c->h_struct_atom0.data(), (ns + 1) * sizeof(long long), cudaMemcpyHostToDevice, st));
  PP_CUDA(cudaMemcpyAsync(c->d_FtoA.p + 9 * (size_t)s0, c->h_FtoA.data() + 9 * (size_t)s0, 9 * (size_t)n_struct * sizeof(double), cudaMemcpyHostToDevice, st));
  PP_CUDA(cudaMemcpyAsync(c->d_mxyz.p + 3 * (size_t)s0, c->h_mxyz.data() + 3 * (size_t)s0, 3 * (size_t)n_struct * sizeof(int), cudaMemcpyHostToDevice, st));
  PP_CUDA(cudaMemcpyAsync(c->d_target.p + s0, c->h_target.data() + s0, n_struct * sizeof(double), cudaMemcpyHostToDevice, st));
  PP_CUDA(cudaMemcpyAsync(c->d_tag.p + s0, c->h_tag.data() + s0, n_struct * sizeof(int), cudaMemcpyHostToDevice, st));
  PP_CUDA(cudaMemcpyAsync(c->d_frac.p + 3 * (size_t)a0, frac, 3 * (size_t)add_atoms * sizeof(double), cudaMemcpyHostToDevice, st));
  PP_CUDA(cudaMemcpyAsync(c->d_elem.p + a0, elem, add_atoms * sizeof(int), cudaMemcpyHostToDevice, st));
  PP_CUDA(cudaMemcpyAsync(c->d_atom_struct.p + a0, c->h_atom_struct.data() + a0, add_atoms * sizeof(int), cudaMemcpyHostToDevice, st));
  c->n_struct = ns;
  c->n_atoms = na;
  c->tiles_dirty = true;
  c->g32_dirty = true;
  c->gn_dirty = true;
  // per-element row lists (structure-major)
  for (int e = 0; e < S.n_elem; e++) {
    size_t before = c->h_rows[e].size();
    for (long long i = 0; i < add_atoms; i++) if (elem[i] == e) c->h_rows[e].push_back((int)(a0 + i));
    size_t after = c->h_rows[e].size();
    if (c->d_rows[e].ensure(std::max<size_t>(after, 1), before, st)) return 1;
    if (after > before)
      PP_CUDA(cudaMemcpyAsync(c->d_rows[e].p + before, c->h_rows[e].data() + before, (after - before) * sizeof(int), cudaMemcpyHostToDevice, st));
  }
  // descriptors of the new atoms
  PeriodicArgs a = periodic_args(c);
  a.n_atoms = add_atoms;
  a.atom_begin = a0;
  if (tick(c, 0)) return 1;
  if (ppk_desc_periodic(S, a, &c->desc_cap, st, &c->launches)) return 1;
  if (tick(c, 1)) return 1;
  PP_CUDA(cudaStreamSynchronize(st));
  return tock(c, 0, 1, 0);
}

// Structure::init_G called again on every system after the rows changed (Potential.cpp:806-808): the positions,
// cells and element lists are still resident, so this is one descriptor launch over all atoms.
extern "C" int pp_recompute_descriptors(pp_context* c)
{
  PP_REQUIRE(c && c->spec_set, "pp_recompute_descriptors: call pp_set_descriptors first");
  if (c->n_atoms == 0) return 0;
  PP_CUDA(cudaSetDevice(c->device));
  PeriodicArgs a = periodic_args(c);
  a.n_atoms = c->n_atoms;
  a.atom_begin = 0;
  c->g32_dirty = true;
  c->gn_dirty = true;
  if (tick(c, 0)) return 1;
  if (ppk_desc_periodic(c->spec, a, &c->desc_cap, c->stream, &c->launches)) return 1;
  if (tick(c, 1)) return 1;
  PP_CUDA(cudaStreamSynchronize(c->stream));
  return tock(c, 0, 1, 0);
}

extern "C" int pp_clear_structures(pp_context* c)
{
  PP_REQUIRE(c, "null context");
  c->n_struct = 0;
  c->n_atoms = 0;
  c->tiles_dirty = true;
  c->h_struct_atom0.assign(1, 0);
  c->h_FtoA.clear(); c->h_target.clear(); c->h_mxyz.clear(); c->h_tag.clear(); c->h_elem.clear();
  c->h_atom_struct.clear(); c->h_frac.clear();
  for (int e = 0; e < PP_MAX_ELEM; e++) c->h_rows[e].clear();
  return 0;
}

extern "C" int pp_num_structures(pp_context* c) { return c ? c->n_struct : -1; }
extern "C" long long pp_num_atoms(pp_context* c) { return c ? c->n_atoms : -1; }

extern "C" int pp_get_descriptors(pp_context* c, int s, double* G_out)
{
  PP_REQUIRE(c && s >= 0 && s < c->n_struct, "pp_get_descriptors: structure %d out of range", s);
  PP_CUDA(cudaSetDevice(c->device));
  const int NG = c->spec.NG;
  const long long a0 = c->h_struct_atom0[s];
  const int n = (int)(c->h_struct_atom0[s + 1] - a0);
  std::vector<double> tmp((size_t)n * NG);
  PP_CUDA(cudaMemcpyAsync(tmp.data(), c->d_G.p + (size_t)a0 * NG, tmp.size() * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  PP_CUDA(cudaStreamSynchronize(c->stream));
  for (int i = 0; i < n; i++)
    for (int g = 0; g < NG; g++) G_out[(size_t)g * n + i] = tmp[(size_t)i * NG + g];
  return 0;
}

extern "C" int pp_get_neighbors(pp_context* c, int s, int* neigh_offset, int* neigh_j, int* image_offset, int* image_t,
                                int* n_pairs, int* n_images)
{
  PP_REQUIRE(c && s >= 0 && s < c->n_struct, "pp_get_neighbors: structure %d out of range", s);
  PP_CUDA(cudaSetDevice(c->device));
  const long long a0 = c->h_struct_atom0[s];
  const int n = (int)(c->h_struct_atom0[s + 1] - a0);
  cudaStream_t st = c->stream;
  DevBuf<int> d_counts, d_offsets, d_j, d_t;
  if (d_counts.ensure(n, 0, st) || d_offsets.ensure(n + 1, 0, st)) return 1;
  PeriodicArgs a = periodic_args(c);
  int cap = c->desc_cap;
  std::vector<int> counts(n), offsets(n + 1, 0);
  int rc = 1;
  do {
    PP_CUDA(cudaMemsetAsync(a.overflow, 0, sizeof(int), st));
    if (ppk_neighbors_periodic(c->spec, a, s, n, 0, d_counts.p, nullptr, nullptr, nullptr, cap, st, &c->launches)) break;
    if (cudaMemcpyAsync(counts.data(), d_counts.p, n * sizeof(int), cudaMemcpyDeviceToHost, st) != cudaSuccess) break;
    if (cudaStreamSynchronize(st) != cudaSuccess) break;
    int mx = 0;
    for (int i = 0; i < n; i++) { offsets[i + 1] = offsets[i] + counts[i]; mx = std::max(mx, counts[i]); }
    while (cap < mx) cap *= 2;
    if (cap > ppk_max_cap(c->spec, false)) { pp_set_error("pp_get_neighbors: too many neighbour images for shared memory"); break; }
    const int total = offsets[n];
    if (d_j.ensure(std::max(total, 1), 0, st) || d_t.ensure(std::max(total, 1), 0, st)) break;
    if (cudaMemcpyAsync(d_offsets.p, offsets.data(), (n + 1) * sizeof(int), cudaMemcpyHostToDevice, st) != cudaSuccess) break;
    if (ppk_neighbors_periodic(c->spec, a, s, n, 1, d_counts.p, d_offsets.p, d_j.p, d_t.p, cap, st, &c->launches)) break;
    std::vector<int> hj(total), ht(total);
    if (total) {
      if (cudaMemcpyAsync(hj.data(), d_j.p, total * sizeof(int), cudaMemcpyDeviceToHost, st) != cudaSuccess) break;
      if (cudaMemcpyAsync(ht.data(), d_t.p, total * sizeof(int), cudaMemcpyDeviceToHost, st) != cudaSuccess) break;
    }
    if (cudaStreamSynchronize(st) != cudaSuccess) break;
    // (j, image) pairs -> unique-j list + per-pair image lists, the layout of Structure::firstneigh / trans_indices
    int np = 0, ni = 0;
    for (int i = 0; i < n; i++) {
      if (neigh_offset) neigh_offset[i] = np;
      int last_j = -1;
      for (int p = offsets[i]; p < offsets[i + 1]; p++) {
        if (hj[p] != last_j) {
          if (neigh_j) { neigh_j[np] = hj[p]; image_offset[np] = ni; }
          np++;
          last_j = hj[p];
        }
        if (image_t) image_t[ni] = ht[p];
        ni++;
      }
    }
    if (neigh_offset) neigh_offset[n] = np;
    if (image_offset && neigh_j) image_offset[np] = ni;
    if (n_pairs) *n_pairs = np;
    if (n_images) *n_images = ni;
    rc = 0;
  } while (0);
  d_counts.release(); d_offsets.release(); d_j.release(); d_t.release();
  if (rc && !g_err[0]) pp_set_error("pp_get_neighbors: CUDA failure (%s)", cudaGetErrorString(cudaGetLastError()));
  return rc;
}

// ---------------------------------------------------------------------------------------------------------
// Potential::Precondition, input part (Potential.cpp:404-455)
extern "C" int pp_precondition(pp_context* c, int input_precondition)
{
  if (nets_ready(c)) return 1;
  PP_CUDA(cudaSetDevice(c->device));
  const int NG = c->spec.NG;
  const int n_cta = 2 * ppk_mlp_single_ctas();
  cudaStream_t st = c->stream;
  if (c->d_stat_partial.ensure((size_t)n_cta * NG, 0, st) || c->d_stat_sums.ensure(2 * (size_t)NG, 0, st) || c->d_count.ensure(1, 0, st)) return 1;
  for (int e = 0; e < c->spec.n_elem; e++) {
    std::vector<double> mean(NG, 0.0), var(NG, 1.0);
    if (input_precondition) {
      const int M = (int)c->h_rows[e].size();
      PP_CUDA(cudaMemsetAsync(c->d_count.p, 0, sizeof(long long), st));
      if (ppk_column_stats(c->d_G.p, NG, c->d_rows[e].p, M, c->d_atom_struct.p, c->d_tag.p, nullptr, 0, c->d_stat_partial.p, n_cta,
                           c->d_stat_sums.p, c->d_count.p, st, &c->launches)) return 1;
      std::vector<double> buf(NG + 1);
      long long cnt = 0;
      PP_CUDA(cudaMemcpyAsync(buf.data(), c->d_stat_sums.p, NG * sizeof(double), cudaMemcpyDeviceToHost, st));
      PP_CUDA(cudaMemcpyAsync(&cnt, c->d_count.p, sizeof(long long), cudaMemcpyDeviceToHost, st));
      PP_CUDA(cudaStreamSynchronize(st));
      buf[NG] = (double)cnt;
      if (c->comm.allreduce_host(buf.data(), NG + 1, st)) return 1;    // Potential.cpp:419-420
      double count = buf[NG];
      for (int k = 0; k < NG; k++) mean[k] = count != 0 ? buf[k] / count : buf[k];
      PP_CUDA(cudaMemcpyAsync(c->d_stat_sums.p + NG, mean.data(), NG * sizeof(double), cudaMemcpyHostToDevice, st));
      if (ppk_column_stats(c->d_G.p, NG, c->d_rows[e].p, M, c->d_atom_struct.p, c->d_tag.p, c->d_stat_sums.p + NG, 1, c->d_stat_partial.p,
                           n_cta, c->d_stat_sums.p, c->d_count.p, st, &c->launches)) return 1;
      PP_CUDA(cudaMemcpyAsync(buf.data(), c->d_stat_sums.p, NG * sizeof(double), cudaMemcpyDeviceToHost, st));
      PP_CUDA(cudaStreamSynchronize(st));
      if (c->comm.allreduce_host(buf.data(), NG, st)) return 1;        // Potential.cpp:442
      for (int k = 0; k < NG; k++) {
        var[k] = buf[k];
        if (count != 0) {
          var[k] = sqrt(var[k] / count);
          if (var[k] < 1e-5) var[k] = 1e-5;
        }
      }
    }
    if (pp_set_input_scaling(c, e, mean.data(), var.data())) return 1;
  }
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
static MlpArgs mlp_args_train(pp_context* c, int e)
{
  MlpArgs a;
  memset(&a, 0, sizeof(a));
  a.net = c->nets[e].dev;
  const bool gn = c->gn_ready && !c->gn_dirty;      // never a copy that descriptors or scalings have moved away from
  a.G = gn ? c->d_Gn.p : c->d_G.p;
  a.prenorm = gn ? 1 : 0;
  a.rows = c->spec.n_elem == 1 ? nullptr : c->d_rows[e].p;
  a.M = (int)c->h_rows[e].size();
  a.atom_struct = c->d_atom_struct.p;
  a.cstruct = c->d_cstruct.p;
  a.out_scale = c->out_scale;
  a.E_atom = c->d_E_atom.p;
  a.partial = c->d_partial.p;
  a.P_total = c->P_total;
  a.p_off = c->nets[e].p_off;
  a.tanh_table = c->d_tanh_table.p;
  a.dw_tiles = c->nets[e].d_dw_tiles.p;
  a.dw_tiles2 = c->nets[e].has_tiles2 ? c->nets[e].d_dw_tiles2.p : nullptr;
  a.scratch2 = c->nets[e].has_tiles2 ? c->d_scratch2.p : nullptr;
  a.n_tiles = c->n_fused_tiles;
  a.tile_struct0 = c->d_tile_struct0.p;
  a.struct_atom0 = c->d_struct_atom0.p;
  a.target = c->d_target.p;
  a.tag = c->d_tag.p;
  a.out_mean = c->out_mean;
  a.E_struct = c->d_E_struct.p;
  return a;
}

// Single-element systems whose structures each fit one MLP tile take the fused single-pass epoch: tiles are
// runs of whole consecutive structures, packed greedily up to the tile size.
static bool fp32_epoch(pp_context* c)
{
  return c->fp_mode == 32 && c->spec.n_elem == 1 && c->nets[0].set && c->nets[0].has_tiles32 && getenv("PP_FP32_OFF") == nullptr;
}

static int prepare_fused_tiles(pp_context* c)
{
  int TM = 0;
  if (c->spec.n_elem == 1 && c->nets[0].set) {
    TM = c->nets[0].tile_atoms;
    if (fp32_epoch(c)) {
      // the tcgen05 kernel works on 128-atom tiles; a structure of 129 .. 256 atoms sends the system to the mma.sync kernel
      int largest = 0;
      for (int s = 0; s < c->n_struct; s++) largest = std::max(largest, (int)(c->h_struct_atom0[s + 1] - c->h_struct_atom0[s]));
      c->fp32_umma = ppk_mlp32u_ok(c->nets[0].dev) && largest <= ppk_mlp32u_tile_atoms();
      TM = c->fp32_umma ? ppk_mlp32u_tile_atoms() : ppk_mlp32_tile_atoms();
    }
  }
  if (!c->tiles_dirty && c->tiles_tm == TM) return 0;
  c->tiles_dirty = false;
  c->tiles_tm = TM;
  c->fused_ok = false;
  c->n_fused_tiles = 0;
  if (c->spec.n_elem != 1 || !c->nets[0].set || c->n_struct == 0) return 0;
  std::vector<int> t0;
  t0.push_back(0);
  int used = 0;
  for (int s = 0; s < c->n_struct; s++) {
    const int n = (int)(c->h_struct_atom0[s + 1] - c->h_struct_atom0[s]);
    if (n > TM) return 0;
    if (used + n > TM) { t0.push_back(s); used = 0; }
    used += n;
  }
  t0.push_back(c->n_struct);
  if (c->d_tile_struct0.ensure(t0.size(), 0, c->stream)) return 1;
  PP_CUDA(cudaMemcpyAsync(c->d_tile_struct0.p, t0.data(), t0.size() * sizeof(int), cudaMemcpyHostToDevice, c->stream));
  PP_CUDA(cudaStreamSynchronize(c->stream));
  c->n_fused_tiles = (int)t0.size() - 1;
  c->fused_ok = true;
  return 0;
}

// every element's network can run the two-CTA kernels (then all of them do: the partial-gradient rows must agree)
static bool all_two_cta(pp_context* c)
{
  for (int e = 0; e < c->spec.n_elem; e++)
    if (!c->nets[e].has_tiles2 || !ppk_mlp_two_cta_ok(c->nets[e].dev)) return false;
  return true;
}

// every element's network can run the two-phase kernels that keep the activations in HBM between the phases
static bool all_stored(pp_context* c)
{
  for (int e = 0; e < c->spec.n_elem; e++)
    if (!c->nets[e].has_tiles3 || !ppk_mlpw_ok(c->nets[e].dev)) return false;
  return true;
}

static MlpwStore store_of(pp_context* c, int e, bool with_store)
{
  MlpwStore st;
  st.H = with_store ? c->d_hstore[e].p : nullptr;
  st.T = with_store ? c->d_tstore[e].p : nullptr;
  return st;
}

static int struct_errors_all(pp_context* c)
{
  StructArgs sa;
  sa.n_struct = c->n_struct;
  sa.struct_atom0 = c->d_struct_atom0.p;
  sa.E_atom = c->d_E_atom.p;
  sa.target = c->d_target.p;
  sa.tag = c->d_tag.p;
  sa.out_mean = c->out_mean;
  sa.out_scale = c->out_scale;
  sa.E_struct = c->d_E_struct.p;
  sa.cstruct = c->d_cstruct.p;
  sa.result_tail = c->d_result.p + c->P_total;
  sa.scratch = c->d_err_scratch.p;
  sa.ticket = c->d_err_ticket.p;
  return ppk_struct_errors(sa, c->stream, &c->launches);
}

// forward pass of every element + per-structure energies and errors.  stored: the kernels of pp_mlpw.cu, which also
// keep H_l / T'_l for the backward pass when with_store is set.
// Bring the normalised copy of the descriptors up to date (or switch it off: PP_MLP_NO_PRENORM=1, or no memory for it).
static int ensure_normalised(pp_context* c)
{
  c->gn_ready = false;
  if (getenv("PP_MLP_NO_PRENORM") != nullptr || c->n_atoms == 0) return 0;
  const int NG = c->spec.NG, ne = c->spec.n_elem;
  if (c->gn_dirty) {
    double* p = nullptr;
    const size_t need = (size_t)c->n_atoms * NG;
    if (c->d_Gn.cap < need) {               // a failed allocation only costs the optimisation
      c->d_Gn.release();
      if (cudaMalloc((void**)&p, need * sizeof(double)) != cudaSuccess) { cudaGetLastError(); return 0; }
      c->d_Gn.p = p; c->d_Gn.cap = need;
    }
    std::vector<double> ms(2 * (size_t)ne * NG);
    for (int e = 0; e < ne; e++)
      for (int k = 0; k < NG; k++) {
        ms[(size_t)e * NG + k] = c->nets[e].mean[k];
        ms[(size_t)(ne + e) * NG + k] = 1.0 / c->nets[e].scale[k];
      }
    if (c->d_msn.ensure(ms.size(), 0, c->stream)) return 1;
    PP_CUDA(cudaMemcpyAsync(c->d_msn.p, ms.data(), ms.size() * sizeof(double), cudaMemcpyHostToDevice, c->stream));
    PP_CUDA(cudaStreamSynchronize(c->stream));      // `ms` is a local
    if (ppk_g_normalise(c->d_G.p, NG, c->n_atoms, ne == 1 ? nullptr : c->d_elem.p, c->d_msn.p, c->d_msn.p + (size_t)ne * NG, c->d_Gn.p, c->stream,
                        &c->launches))
      return 1;
    c->gn_dirty = false;
  }
  c->gn_ready = true;
  return 0;
}

// the two-CTA kernels' own stored form (epoch form 5): FORWARD writes, TRAIN reads the element's activation store
static void with_store2(pp_context* c, int e, MlpArgs& a)
{
  a.hstore = c->d_hstore[e].p;
  a.tstore = c->d_tstore[e].p;
  a.store_w = ppk_mlpw_store_width(c->nets[e].dev);
}

static int forward_all(pp_context* c, bool stored, bool with_store, bool store2 = false)
{
  const bool two = !stored && all_two_cta(c);
  for (int e = 0; e < c->spec.n_elem; e++) {
    MlpArgs a = mlp_args_train(c, e);
    if (stored) {
      if (ppk_mlpw_forward(a, store_of(c, e, with_store), c->stream, &c->launches)) return 1;
    } else {
      int n_cta = 0;
      if (store2) with_store2(c, e, a);
      if (ppk_mlp_launch2(PP_MLP_FORWARD, two, a, c->stream, &c->launches, &n_cta)) return 1;
    }
  }
  return struct_errors_all(c);
}

static int epoch_buffers(pp_context* c)
{
  const int n_cta = ppk_mlp_max_ctas();
  if (c->d_partial.ensure((size_t)n_cta * (c->P_total + 3), 0, c->stream)) return 1;
  if (c->d_result.ensure((size_t)c->P_total + 3, 0, c->stream)) return 1;
  if (c->d_err_scratch.ensure(3 * (size_t)ppk_struct_errors_blocks(std::max(c->n_struct, 1)), 0, c->stream)) return 1;
  size_t sc2 = 0;
  for (int e = 0; e < c->spec.n_elem; e++)
    if (c->nets[e].has_tiles2) sc2 = std::max(sc2, ppk_mlp2_scratch_doubles(c->nets[e].dev));
  if (sc2 && c->d_scratch2.ensure(sc2, 0, c->stream)) return 1;
  if (!c->d_err_ticket.p) {
    if (c->d_err_ticket.ensure(1, 0, c->stream)) return 1;
    PP_CUDA(cudaMemsetAsync(c->d_err_ticket.p, 0, sizeof(unsigned), c->stream));
  }
  return 0;
}

// Which form the epoch takes: 1 / 2 = single pass (single-element system whose structures fit a tile), 3 = two phases
// with the activations kept in HBM in between (several elements, or a network too large for 64-atom tiles in the
// single-pass kernels), 0 = two phases with the forward pass recomputed (anything else).
static int choose_epoch_form(pp_context* c)
{
  // networks of up to five n-tiles and two hidden layers are quicker through the two-CTA kernels' FORWARD / TRAIN modes
  // (measured on C3); PP_MLP_FORCE_STORE=1 sends every eligible system through the stored form
  const bool stored = all_stored(c) && (!all_two_cta(c) || getenv("PP_MLP_FORCE_STORE") != nullptr);
  if (c->fused_ok) {
    const NetHost& n = c->nets[0];
    const bool two = n.has_tiles2 && ppk_mlp_two_cta_ok(n.dev);
    if (two) return 2;
    if (stored && n.tile_atoms < 64) return 3;
    return 1;
  }
  if (stored) return 3;
  // small networks (the two-CTA kernels): their FORWARD launch keeps H_l and T'_l in HBM and the TRAIN launch reads them
  // back instead of running the forward pass again; PP_MLP_NO_STORE=1 keeps the recomputing form
  if (all_two_cta(c) && ppk_mlp2_available() && getenv("PP_MLP_NO_STORE") == nullptr) return 5;
  return 0;
}

extern "C" int pp_epoch_device(pp_context* c)
{
  if (nets_ready(c)) return 1;
  // a single rank without structures has nothing to train on; one rank of several (more ranks than systems, which
  // the reference tolerates, Setup.cpp:582) contributes zeros and still joins the reduction
  PP_REQUIRE(c->n_struct > 0 || c->comm.nranks > 1, "pp_epoch: no structures");
  PP_CUDA(cudaSetDevice(c->device));
  if (epoch_buffers(c)) return 1;
  if (prepare_fused_tiles(c)) return 1;
  if (!(fp32_epoch(c) && c->fused_ok) && ensure_normalised(c)) return 1;      // (the FP32 epoch reads its own FP32 copy)
  if (tick(c, 0)) return 1;
  const int stride = c->P_total + 3;
  // what the reduction at the end sums: the partial rows' columns [0, red_cols), then d_result's own [red_cols, stride)
  int red_rows = 0, red_cols = 0;
  if (c->n_struct == 0) {
    if (tick(c, 1)) return 1;
    PP_CUDA(cudaMemsetAsync(c->d_result.p, 0, (size_t)stride * sizeof(double), c->stream));
    c->epoch_form = -1;
  } else {
    const int form = (fp32_epoch(c) && c->fused_ok) ? 4 : choose_epoch_form(c);
    c->epoch_form = form;
    if (form == 4) {
      // FP32 mode: one pass on 3xTF32 products over the normalised FP32 copy of the descriptors (refreshed when the
      // descriptors or the input scaling changed)
      const int NG = c->spec.NG, NGp = ppk_mlp32_pitch(c->nets[0].dev);
      if (c->g32_dirty) {
        if (c->d_G32.ensure((size_t)c->n_atoms * NGp, 0, c->stream) || c->d_ms32.ensure(2 * (size_t)NG, 0, c->stream)) return 1;
        PP_CUDA(cudaMemcpyAsync(c->d_ms32.p, c->nets[0].mean.data(), NG * sizeof(double), cudaMemcpyHostToDevice, c->stream));
        PP_CUDA(cudaMemcpyAsync(c->d_ms32.p + NG, c->nets[0].scale.data(), NG * sizeof(double), cudaMemcpyHostToDevice, c->stream));
        if (ppk_g_to_f32(c->d_G.p, NG, NGp, c->n_atoms, nullptr, c->d_ms32.p, c->d_ms32.p + NG, c->d_G32.p, c->stream, &c->launches)) return 1;
        c->g32_dirty = false;
      }
      if (tick(c, 1)) return 1;
      MlpArgs a = mlp_args_train(c, 0);
      a.dw_tiles2 = c->nets[0].d_dw_tiles32.p;
      int n_cta = 0;
      if (c->fp32_umma) {          // forward and delta products on tcgen05.mma (pp_mlp32u.cu)
        if (ppk_mlp32u_launch_fused(a, c->d_G32.p, c->n_atoms, c->stream, &c->launches, &n_cta)) return 1;
        c->fp32_umma_ran = true;
      } else if (ppk_mlp32_launch_fused(a, c->d_G32.p, c->n_atoms, c->stream, &c->launches, &n_cta)) return 1;
      red_rows = n_cta; red_cols = stride;
    } else if (form == 1 || form == 2) {
      // one pass: forward, per-structure error and backward in the same kernel; loss terms ride in the partials' tail
      if (tick(c, 1)) return 1;
      MlpArgs a = mlp_args_train(c, 0);
      int n_cta = 0;
      if (ppk_mlp_launch_fused(a, c->stream, &c->launches, &n_cta)) return 1;
      c->epoch_form = (n_cta > ppk_mlp_single_ctas()) ? 2 : 1;
      red_rows = n_cta; red_cols = stride;
    } else {
      const bool stored = (form == 3), store2 = (form == 5);
      if (stored || store2) {
        for (int e = 0; e < c->spec.n_elem; e++) {
          const size_t need = c->h_rows[e].size() * (size_t)ppk_mlpw_store_width(c->nets[e].dev);
          if (c->d_hstore[e].ensure(std::max<size_t>(need, 2), 0, c->stream) || c->d_tstore[e].ensure(std::max<size_t>(need, 2), 0, c->stream)) return 1;
        }
      }
      if (forward_all(c, stored, true, store2)) return 1;
      if (tick(c, 1)) return 1;
      const bool two = !stored && all_two_cta(c);
      int rows_e[PP_MAX_ELEM] = {0, 0, 0, 0}, n_max = 0;
      for (int e = 0; e < c->spec.n_elem; e++) {
        MlpArgs a = mlp_args_train(c, e);
        if (a.M == 0) continue;                 // this rank holds no atom of the element: its columns are zeroed below
        if (stored) {
          a.dw_tiles2 = c->nets[e].d_dw_tiles3.p;
          if (ppk_mlpw_backward(a, store_of(c, e, true), c->stream, &c->launches, &rows_e[e])) return 1;
        } else {
          if (store2) with_store2(c, e, a);
          if (ppk_mlp_launch2(PP_MLP_TRAIN, two, a, c->stream, &c->launches, &rows_e[e])) return 1;
        }
        n_max = std::max(n_max, rows_e[e]);
      }
      if (n_max == 0) n_max = 1;
      // every element's column block must hold n_max rows: zero the rows its launch did not write
      for (int e = 0; e < c->spec.n_elem; e++) {
        if (rows_e[e] >= n_max || c->nets[e].P == 0) continue;
        PP_CUDA(cudaMemset2DAsync(c->d_partial.p + (size_t)rows_e[e] * stride + c->nets[e].p_off, (size_t)stride * sizeof(double), 0,
                                  (size_t)c->nets[e].P * sizeof(double), (size_t)(n_max - rows_e[e]), c->stream));
      }
      red_rows = n_max; red_cols = c->P_total;     // the loss terms are in d_result's tail already (struct_errors_kernel)
    }
  }
  if (tick(c, 2)) return 1;
  // fixed-order sum of the partial rows and, with several ranks, the sum over ranks (Optimizer.cpp:887-888, 969-970)
  if (c->comm.reduce_sum(c->d_partial.p, red_rows, stride, red_cols, stride, c->d_result.p, c->stream, &c->launches)) return 1;
  if (tick(c, 3)) return 1;
  return 0;
}

extern "C" int pp_epoch_form(pp_context* c) { return c ? c->epoch_form : -1; }

extern "C" int pp_epoch_result_dev(pp_context* c, double** result_dev, int* n)
{
  PP_REQUIRE(c && result_dev && n, "pp_epoch_result_dev: bad arguments");
  PP_CUDA(cudaSetDevice(c->device));
  if (epoch_buffers(c)) return 1;
  *result_dev = c->d_result.p;
  *n = c->P_total + 3;
  return 0;
}

static double now_us()
{
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return ts.tv_sec * 1e6 + ts.tv_nsec * 1e-3;
}
static double g_trace[4] = {0, 0, 0, 0};
static long g_trace_n = 0;

extern "C" int pp_epoch(pp_context* c, const double* params_in, double* grad_out, double* sse, double* vsse, long long* nval)
{
  const double t0 = now_us();
  if (nets_ready(c)) return 1;
  PP_CUDA(cudaSetDevice(c->device));
  // pinned staging: [P_total + 3] results, then [P_total] parameters (a DMA each way whatever the caller's memory is)
  const size_t need = 2 * (size_t)c->P_total + 3;
  if (c->h_result_cap < need) {
    if (c->h_result) cudaFreeHost(c->h_result);
    c->h_result = nullptr;
    c->h_result_cap = 0;
    PP_CUDA(cudaMallocHost((void**)&c->h_result, need * sizeof(double)));
    c->h_result_cap = need;
  }
  if (params_in) {
    c->paramsT_dirty = true;
    double* stage = c->h_result + c->P_total + 3;
    for (int e = 0; e < c->spec.n_elem; e++) {
      NetHost& n = c->nets[e];
      std::copy(params_in + n.p_off, params_in + n.p_off + n.P, n.params.begin());
    }
    std::copy(params_in, params_in + c->P_total, stage);
    PP_CUDA(cudaMemcpyAsync(c->d_params.p, stage, c->P_total * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  }
  if (pp_epoch_device(c)) return 1;
  PP_CUDA(cudaMemcpyAsync(c->h_result, c->d_result.p, ((size_t)c->P_total + 3) * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  const double t1 = now_us();
  PP_CUDA(cudaStreamSynchronize(c->stream));
  if (c->fp32_umma_ran) {
    c->fp32_umma_ran = false;
    PP_REQUIRE(!ppk_mlp32u_failed(c->stream), "pp_epoch: a tensor-core batch of the FP32 kernel did not complete (internal error)");
  }
  const double t2 = now_us();
  c->epoch_phases_pending = true;
  const double t3 = now_us();
  if (getenv("PP_EPOCH_TRACE")) {
    g_trace[0] += t1 - t0; g_trace[1] += t2 - t1; g_trace[2] += t3 - t2; g_trace_n++;
    if (g_trace_n % 20 == 0)
      fprintf(stderr, "pp_epoch trace: enqueue %.1f us, wait %.1f us, events %.1f us (mean of %ld)\n", g_trace[0] / g_trace_n,
              g_trace[1] / g_trace_n, g_trace[2] / g_trace_n, g_trace_n);
  }
  if (grad_out) std::copy(c->h_result, c->h_result + c->P_total, grad_out);
  if (sse) *sse = c->h_result[c->P_total];
  if (vsse) *vsse = c->h_result[c->P_total + 1];
  if (nval) *nval = (long long)llround(c->h_result[c->P_total + 2]);
  return 0;
}

extern "C" int pp_evaluate(pp_context* c, double* E_out)
{
  if (nets_ready(c)) return 1;
  PP_REQUIRE(c->n_struct > 0, "pp_evaluate: no structures");
  PP_CUDA(cudaSetDevice(c->device));
  if (epoch_buffers(c)) return 1;
  if (ensure_normalised(c)) return 1;
  if (forward_all(c, all_stored(c) && !all_two_cta(c), false)) return 1;
  PP_CUDA(cudaMemcpyAsync(E_out, c->d_E_struct.p, c->n_struct * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  PP_CUDA(cudaStreamSynchronize(c->stream));
  return 0;
}

extern "C" int pp_get_atom_energies(pp_context* c, int s, double* E_atom_out)
{
  PP_REQUIRE(c && s >= 0 && s < c->n_struct, "pp_get_atom_energies: structure %d out of range", s);
  PP_CUDA(cudaSetDevice(c->device));
  const long long a0 = c->h_struct_atom0[s];
  const int n = (int)(c->h_struct_atom0[s + 1] - a0);
  PP_CUDA(cudaMemcpyAsync(E_atom_out, c->d_E_atom.p + a0, n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  PP_CUDA(cudaStreamSynchronize(c->stream));
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
// LAMMPS pair_style nn (pair_nn.cpp:112-188)
extern "C" int pp_md_set_types(pp_context* c, int ntypes, const int* type_to_elem)
{
  PP_REQUIRE(c && c->spec_set, "pp_md_set_types: descriptors not set");
  PP_REQUIRE(ntypes >= 1, "pp_md_set_types: ntypes must be >= 1");
  c->type_to_elem.assign(type_to_elem, type_to_elem + ntypes + 1);
  for (int t = 1; t <= ntypes; t++)
    PP_REQUIRE(c->type_to_elem[t] >= 0 && c->type_to_elem[t] < c->spec.n_elem, "pp_md_set_types: type %d maps to element %d", t, c->type_to_elem[t]);
  return 0;
}

static int md_upload_atoms(pp_context* c, const double* x, const int* type, int nlocal, int nghost)
{
  const int ntotal = nlocal + nghost;
  cudaStream_t st = c->stream;
  const int NG = c->spec.NG;
  if (c->d_x.ensure(3 * (size_t)ntotal, 0, st) || c->d_f.ensure(3 * (size_t)ntotal, 0, st) || c->d_md_elem.ensure(ntotal, 0, st) ||
      c->d_Gmd.ensure((size_t)std::max(nlocal, 1) * NG, 0, st) || c->d_dEdG.ensure((size_t)std::max(nlocal, 1) * NG, 0, st) ||
      c->d_eatom.ensure(std::max(nlocal, 1), 0, st))
    return 1;
  c->h_md_elem.resize(ntotal);
  const int ntypes = (int)c->type_to_elem.size() - 1;
  for (int i = 0; i < ntotal; i++) {
    PP_REQUIRE(type[i] >= 1 && type[i] <= ntypes, "pp_md: atom %d has type %d outside 1..%d", i, type[i], ntypes);
    c->h_md_elem[i] = c->type_to_elem[type[i]];
  }
  PP_CUDA(cudaMemcpyAsync(c->d_x.p, x, 3 * (size_t)ntotal * sizeof(double), cudaMemcpyHostToDevice, st));
  PP_CUDA(cudaMemcpyAsync(c->d_md_elem.p, c->h_md_elem.data(), ntotal * sizeof(int), cudaMemcpyHostToDevice, st));
  // per-element lists of local atoms
  for (int e = 0; e < c->spec.n_elem; e++) {
    std::vector<int> rows;
    for (int i = 0; i < nlocal; i++) if (c->h_md_elem[i] == e) rows.push_back(i);
    c->md_rows_n[e] = (int)rows.size();
    if (c->d_md_rows[e].ensure(std::max<size_t>(rows.size(), 1), 0, st)) return 1;
    if (!rows.empty()) {
      PP_CUDA(cudaMemcpyAsync(c->d_md_rows[e].p, rows.data(), rows.size() * sizeof(int), cudaMemcpyHostToDevice, st));
      PP_CUDA(cudaStreamSynchronize(st));   // rows is a local
    }
  }
  c->md_nlocal = nlocal;
  c->md_ntotal = ntotal;
  return 0;
}

// transposed copy of every weight layer for the fused MD kernel's forward products (host transposition: P <= ~50k)
static int md_refresh_transposed(pp_context* c)
{
  if (!c->paramsT_dirty) return 0;
  std::vector<double> t((size_t)std::max(c->P_total, 1), 0.0);
  for (int e = 0; e < c->spec.n_elem; e++) {
    const NetHost& n = c->nets[e];
    for (int l = 0; l < n.n_layers; l++) {
      const int n_in = n.sizes[l], n_out = n.sizes[l + 1];
      const double* src = n.params.data() + n.dev.w_off[l];
      double* dst = t.data() + n.p_off + n.dev.w_off[l];
      for (int r = 0; r < n_out; r++)
        for (int k = 0; k <= n_in; k++) dst[(size_t)k * n_out + r] = src[(size_t)r * (n_in + 1) + k];
    }
  }
  if (c->d_paramsT.ensure(t.size(), 0, c->stream)) return 1;
  PP_CUDA(cudaMemcpyAsync(c->d_paramsT.p, t.data(), t.size() * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  PP_CUDA(cudaStreamSynchronize(c->stream));      // t is a local
  c->paramsT_dirty = false;
  return 0;
}

// The single-kernel MD step is opt-in (PP_MD_FUSED=1): measured on the 1 M-atom box it is SLOWER than the three kernels
// (46.0 ms against 34.4 ms; 40.3 ms even with the network left out) although it builds the neighbour shell once -- the
// shell set-up it saves is only 2.0 + 3.7 ms of the 12.0 + 21.3 ms the two pair-loop kernels take, and the combined
// kernel pays for it with spills at the 128-register cap and a code footprint that no longer fits the instruction cache
// with four warps per scheduler at different places in it; the per-atom network costs 5.7 ms in-warp against 1.06 ms
// as a DMMA kernel over all atoms.  Kept because it is exact (tests) and documents the measurement.
static bool md_fused(pp_context* c)
{
  if (getenv("PP_MD_FUSED") == nullptr) return false;
  for (int e = 0; e < c->spec.n_elem; e++)
    if (!ppk_md_fused_ok(c->nets[e].dev)) return false;
  return true;
}

static int md_run(pp_context* c, const double* x_dev, double* f_dev, bool want_G = false, bool want_dEdG = false)
{
  cudaStream_t st = c->stream;
  const int nlocal = c->md_nlocal;
  if (md_fused(c)) {
    // one launch: descriptors, the atom's network and the force chain rule per warp; G and dE/dG stay on the SM
    if (md_refresh_transposed(c)) return 1;
    MdArgs a;
    memset(&a, 0, sizeof(a));
    a.spec = c->d_spec;
    a.nlocal = nlocal;
    a.x = x_dev;
    a.elem = c->d_md_elem.p;
    a.neigh_offsets = c->d_neigh_off.p;
    a.neigh = c->d_neigh.p;
    a.f = f_dev;
    a.overflow = c->d_overflow.p;
    MdNets nets;
    memset(&nets, 0, sizeof(nets));
    for (int e = 0; e < c->spec.n_elem; e++) { nets.net[e] = c->nets[e].dev; nets.WT[e] = c->d_paramsT.p + c->nets[e].p_off; }
    nets.out_scale = c->out_scale;
    nets.tanh_table = c->d_tanh_table.p;
    nets.eatom = c->d_eatom.p;
    nets.G_store = want_G ? c->d_Gmd.p : nullptr;
    nets.dEdG_store = want_dEdG ? c->d_dEdG.p : nullptr;
    if (tick(c, 0)) return 1;
    if (tick(c, 1)) return 1;
    if (tick(c, 2)) return 1;
    if (ppk_md_fused(c->spec, a, nets, &c->md_cap, st, &c->launches)) return 1;
    if (tick(c, 3)) return 1;
    c->md_fused_last = true;
    return 0;
  }
  c->md_fused_last = false;
  MdArgs a;
  memset(&a, 0, sizeof(a));
  a.spec = c->d_spec;
  a.nlocal = nlocal;
  a.x = x_dev;
  a.elem = c->d_md_elem.p;
  a.neigh_offsets = c->d_neigh_off.p;
  a.neigh = c->d_neigh.p;
  a.G = c->d_Gmd.p;
  a.dEdG = c->d_dEdG.p;
  a.f = f_dev;
  a.overflow = c->d_overflow.p;
  if (tick(c, 0)) return 1;
  if (ppk_md_desc(c->spec, a, &c->md_cap, st, &c->launches)) return 1;
  if (tick(c, 1)) return 1;
  for (int e = 0; e < c->spec.n_elem; e++) {
    MlpArgs m;
    memset(&m, 0, sizeof(m));
    m.net = c->nets[e].dev;
    m.G = c->d_Gmd.p;
    m.rows = c->spec.n_elem == 1 ? nullptr : c->d_md_rows[e].p;
    m.M = c->md_rows_n[e];
    m.out_scale = c->out_scale;
    m.E_atom = c->d_eatom.p;
    m.dEdG = c->d_dEdG.p;
    m.tanh_table = c->d_tanh_table.p;
    m.dw_tiles = c->nets[e].d_dw_tiles.p;
    if (ppk_mlp_launch(PP_MLP_INPUT_GRAD, m, st, &c->launches)) return 1;
  }
  if (tick(c, 2)) return 1;
  if (ppk_md_force(c->spec, a, c->md_cap, st, &c->launches)) return 1;
  if (tick(c, 3)) return 1;
  return 0;
}

extern "C" int pp_md_compute(pp_context* c, const double* x, const int* type, int nlocal, int nghost,
                             const long long* neigh_offsets, const int* neigh_flat,
                             double* f_inout, double* eatom_out, double* etotal_out, double* G_out, double* dEdG_out)
{
  if (nets_ready(c)) return 1;
  PP_REQUIRE(!c->type_to_elem.empty(), "pp_md_compute: call pp_md_set_types first");
  PP_REQUIRE(nlocal >= 0 && nghost >= 0 && x && type && neigh_offsets, "pp_md_compute: bad arguments");
  PP_CUDA(cudaSetDevice(c->device));
  cudaStream_t st = c->stream;
  const int ntotal = nlocal + nghost;
  const int NG = c->spec.NG;
  if (md_upload_atoms(c, x, type, nlocal, nghost)) return 1;
  const long long nn = neigh_offsets[nlocal];
  for (long long p = 0; p < nn; p++) PP_REQUIRE(neigh_flat[p] >= 0 && neigh_flat[p] < ntotal, "pp_md_compute: neighbour index %d out of range", neigh_flat[p]);
  if (c->d_neigh_off.ensure(nlocal + 1, 0, st) || c->d_neigh.ensure(std::max<long long>(nn, 1), 0, st)) return 1;
  PP_CUDA(cudaMemcpyAsync(c->d_neigh_off.p, neigh_offsets, (nlocal + 1) * sizeof(long long), cudaMemcpyHostToDevice, st));
  if (nn) PP_CUDA(cudaMemcpyAsync(c->d_neigh.p, neigh_flat, nn * sizeof(int), cudaMemcpyHostToDevice, st));
  c->md_nneigh = nn;
  PP_CUDA(cudaMemsetAsync(c->d_f.p, 0, 3 * (size_t)ntotal * sizeof(double), st));
  c->md_list_resident = false;
  if (md_run(c, c->d_x.p, c->d_f.p, G_out != nullptr, dEdG_out != nullptr)) return 1;
  std::vector<double> f(3 * (size_t)ntotal), ea(std::max(nlocal, 1));
  PP_CUDA(cudaMemcpyAsync(f.data(), c->d_f.p, f.size() * sizeof(double), cudaMemcpyDeviceToHost, st));
  PP_CUDA(cudaMemcpyAsync(ea.data(), c->d_eatom.p, nlocal * sizeof(double), cudaMemcpyDeviceToHost, st));
  if (G_out) PP_CUDA(cudaMemcpyAsync(G_out, c->d_Gmd.p, (size_t)nlocal * NG * sizeof(double), cudaMemcpyDeviceToHost, st));
  if (dEdG_out) PP_CUDA(cudaMemcpyAsync(dEdG_out, c->d_dEdG.p, (size_t)nlocal * NG * sizeof(double), cudaMemcpyDeviceToHost, st));
  PP_CUDA(cudaStreamSynchronize(st));
  tock(c, 0, 1, 4); tock(c, 1, 2, 5); tock(c, 2, 3, 6);
  if (f_inout) for (size_t i = 0; i < f.size(); i++) f_inout[i] += f[i];
  double et = 0.0;
  for (int i = 0; i < nlocal; i++) { et += ea[i]; if (eatom_out) eatom_out[i] = ea[i]; }
  if (etotal_out) *etotal_out = et + c->out_mean;      // pair_nn.cpp:175-179
  return 0;
}

// GPU cell list -> device-resident full neighbour list (LAMMPS' test rsq < cutneigh^2 over locals + ghosts)
extern "C" int pp_md_build_neighbors(pp_context* c, const double* x, const int* type, int nlocal, int nghost, double cutneigh)
{
  if (nets_ready(c)) return 1;
  PP_REQUIRE(!c->type_to_elem.empty(), "pp_md_build_neighbors: call pp_md_set_types first");
  PP_REQUIRE(nlocal >= 0 && nghost >= 0 && x && type && cutneigh > 0, "pp_md_build_neighbors: bad arguments");
  PP_REQUIRE(cutneigh >= c->spec.Rcut, "pp_md_build_neighbors: cutneigh %g is smaller than Rcut %g", cutneigh, c->spec.Rcut);
  PP_CUDA(cudaSetDevice(c->device));
  cudaStream_t st = c->stream;
  const int ntotal = nlocal + nghost;
  if (md_upload_atoms(c, x, type, nlocal, nghost)) return 1;
  // bounding box -> cells of edge >= cutneigh
  CellGrid g;
  double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
  for (int i = 0; i < ntotal; i++)
    for (int k = 0; k < 3; k++) {
      double v = x[3 * (size_t)i + k];
      if (v < lo[k]) lo[k] = v;
      if (v > hi[k]) hi[k] = v;
    }
  long long ncells = 1;
  for (int k = 0; k < 3; k++) {
    if (ntotal == 0) { lo[k] = 0; hi[k] = 1; }
    double L = hi[k] - lo[k];
    int n = (int)floor(L / cutneigh);
    if (n < 1) n = 1;
    if (n > 1024) n = 1024;
    g.n[k] = n;
    g.lo[k] = lo[k];
    g.inv[k] = L > 0 ? n / L : 0.0;
    ncells *= n;
  }
  if (c->d_cell_of.ensure(std::max(ntotal, 1), 0, st) || c->d_sorted.ensure(std::max(ntotal, 1), 0, st) ||
      c->d_cell_count.ensure(ncells, 0, st) || c->d_cell_start.ensure(ncells + 1, 0, st) || c->d_counts.ensure(std::max(nlocal, 1), 0, st) ||
      c->d_scan_sums.ensure(std::max<long long>(ncells, nlocal) / 1024 + 8, 0, st) || c->d_scan_total.ensure(1, 0, st) ||
      c->d_neigh_off.ensure(nlocal + 1, 0, st))
    return 1;
  CellWork w;
  w.cell_of = c->d_cell_of.p; w.cell_count = c->d_cell_count.p; w.cell_start = c->d_cell_start.p; w.sorted = c->d_sorted.p;
  w.counts = c->d_counts.p; w.block_sums = c->d_scan_sums.p; w.total = c->d_scan_total.p;
  if (ppk_build_cell_neighbors(c->d_x.p, nlocal, ntotal, g, cutneigh, w, st, &c->launches)) return 1;
  if (ppk_scan_counts(c->d_counts.p, nlocal, c->d_neigh_off.p, w, st, &c->launches)) return 1;
  long long total = 0;
  PP_CUDA(cudaMemcpyAsync(&total, c->d_scan_total.p, sizeof(long long), cudaMemcpyDeviceToHost, st));
  PP_CUDA(cudaStreamSynchronize(st));
  if (c->d_neigh.ensure(std::max<long long>(total, 1), 0, st)) return 1;
  if (ppk_fill_cell_neighbors(c->d_x.p, nlocal, g, cutneigh, w, c->d_neigh_off.p, c->d_neigh.p, st, &c->launches)) return 1;
  PP_CUDA(cudaStreamSynchronize(st));
  c->md_nneigh = total;
  c->md_list_resident = true;
  return 0;
}

extern "C" int pp_md_get_neighbors(pp_context* c, long long* neigh_offsets, int* neigh_flat, long long* n_total)
{
  PP_REQUIRE(c && c->md_list_resident, "pp_md_get_neighbors: no resident list (call pp_md_build_neighbors)");
  PP_CUDA(cudaSetDevice(c->device));
  if (n_total) *n_total = c->md_nneigh;
  if (neigh_offsets) PP_CUDA(cudaMemcpyAsync(neigh_offsets, c->d_neigh_off.p, (c->md_nlocal + 1) * sizeof(long long), cudaMemcpyDeviceToHost, c->stream));
  if (neigh_flat && c->md_nneigh) PP_CUDA(cudaMemcpyAsync(neigh_flat, c->d_neigh.p, c->md_nneigh * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  PP_CUDA(cudaStreamSynchronize(c->stream));
  return 0;
}

static int md_energy_total(pp_context* c, double* out_dev)
{
  if (c->d_sum_partial.ensure(128, 0, c->stream)) return 1;
  return ppk_sum(c->d_eatom.p, c->md_nlocal, c->out_mean, c->d_sum_partial.p, out_dev, c->stream, &c->launches);
}

// One MD step on the resident list with device-resident positions / forces (forces are accumulated into f_dev).
extern "C" int pp_md_step_device(pp_context* c, const double* x_dev, double* f_dev, double* etotal_dev)
{
  PP_REQUIRE(c && c->md_list_resident, "pp_md_step_device: no resident list (call pp_md_build_neighbors)");
  PP_REQUIRE(x_dev && f_dev, "pp_md_step_device: NULL device pointer");
  PP_CUDA(cudaSetDevice(c->device));
  if (md_run(c, x_dev, f_dev)) return 1;
  if (etotal_dev && md_energy_total(c, etotal_dev)) return 1;
  return 0;
}

// One MD step on the resident list: x uploaded, forces downloaded and ADDED into f_inout, energies downloaded.
extern "C" int pp_md_step(pp_context* c, const double* x, double* f_inout, double* eatom_out, double* etotal_out)
{
  PP_REQUIRE(c && c->md_list_resident, "pp_md_step: no resident list (call pp_md_build_neighbors)");
  PP_REQUIRE(x, "pp_md_step: NULL positions");
  PP_CUDA(cudaSetDevice(c->device));
  cudaStream_t st = c->stream;
  const int ntotal = c->md_ntotal, nlocal = c->md_nlocal;
  if (c->d_etotal.ensure(1, 0, st)) return 1;
  PP_CUDA(cudaMemcpyAsync(c->d_x.p, x, 3 * (size_t)ntotal * sizeof(double), cudaMemcpyHostToDevice, st));
  if (f_inout) PP_CUDA(cudaMemcpyAsync(c->d_f.p, f_inout, 3 * (size_t)ntotal * sizeof(double), cudaMemcpyHostToDevice, st));
  else PP_CUDA(cudaMemsetAsync(c->d_f.p, 0, 3 * (size_t)ntotal * sizeof(double), st));
  if (md_run(c, c->d_x.p, c->d_f.p)) return 1;
  if (md_energy_total(c, c->d_etotal.p)) return 1;
  if (f_inout) PP_CUDA(cudaMemcpyAsync(f_inout, c->d_f.p, 3 * (size_t)ntotal * sizeof(double), cudaMemcpyDeviceToHost, st));
  if (eatom_out) PP_CUDA(cudaMemcpyAsync(eatom_out, c->d_eatom.p, nlocal * sizeof(double), cudaMemcpyDeviceToHost, st));
  double et = 0.0;
  PP_CUDA(cudaMemcpyAsync(&et, c->d_etotal.p, sizeof(double), cudaMemcpyDeviceToHost, st));
  PP_CUDA(cudaStreamSynchronize(st));
  tock(c, 0, 1, 4); tock(c, 1, 2, 5); tock(c, 2, 3, 6);
  if (etotal_out) *etotal_out = et;
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
// Reverse communication of ghost forces for a box partitioned over the ranks (what LAMMPS' Comm::reverse_comm does for
// pair_nn, which leaves forces on ghost atoms, pair_nn.cpp:182).
static int build_seglist(pp_context* c, pp_context::SegList& L, int n, const int* src, const int* dst, int n_dst_max, const char* what)
{
  // stable sort of the values by destination atom -> one segment per destination, values in list order inside it
  std::vector<int> order(n);
  for (int k = 0; k < n; k++) {
    PP_REQUIRE(dst[k] >= 0 && dst[k] < n_dst_max, "pp_md_exchange_setup: %s destination %d out of range", what, dst[k]);
    order[k] = k;
  }
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return dst[a] < dst[b]; });
  std::vector<int> seg_off, seg_dst, idx(n);
  for (int k = 0; k < n; k++) {
    const int o = order[k];
    if (k == 0 || dst[o] != dst[order[k - 1]]) { seg_off.push_back(k); seg_dst.push_back(dst[o]); }
    idx[k] = src ? src[o] : o;
  }
  seg_off.push_back(n);
  L.n_seg = (int)seg_dst.size();
  cudaStream_t st = c->stream;
  if (L.seg_off.ensure(seg_off.size(), 0, st) || L.seg_dst.ensure(std::max<size_t>(seg_dst.size(), 1), 0, st) || L.idx.ensure(std::max(n, 1), 0, st)) return 1;
  PP_CUDA(cudaMemcpyAsync(L.seg_off.p, seg_off.data(), seg_off.size() * sizeof(int), cudaMemcpyHostToDevice, st));
  if (L.n_seg) PP_CUDA(cudaMemcpyAsync(L.seg_dst.p, seg_dst.data(), seg_dst.size() * sizeof(int), cudaMemcpyHostToDevice, st));
  if (n) PP_CUDA(cudaMemcpyAsync(L.idx.p, idx.data(), n * sizeof(int), cudaMemcpyHostToDevice, st));
  PP_CUDA(cudaStreamSynchronize(st));      // the host vectors are locals
  return 0;
}

extern "C" int pp_md_exchange_setup(pp_context* c, int n_self, const int* self_src, const int* self_dst, int n_peers, const int* peer_rank,
                                    const int* send_count, const int* send_idx_flat, const int* recv_count, const int* recv_dst_flat)
{
  PP_REQUIRE(c && c->md_ntotal > 0, "pp_md_exchange_setup: upload the atoms first (pp_md_build_neighbors)");
  PP_REQUIRE(n_self >= 0 && n_peers >= 0, "pp_md_exchange_setup: negative count");
  PP_REQUIRE(n_peers == 0 || c->comm.nranks > 1, "pp_md_exchange_setup: peers given but no communicator (pp_comm_init)");
  PP_CUDA(cudaSetDevice(c->device));
  const int nlocal = c->md_nlocal, ntotal = c->md_ntotal;
  for (int k = 0; k < n_self; k++)
    PP_REQUIRE(self_src[k] >= nlocal && self_src[k] < ntotal, "pp_md_exchange_setup: own ghost index %d is not a ghost", self_src[k]);
  if (build_seglist(c, c->md_self, n_self, self_src, self_dst, nlocal, "own-ghost")) return 1;
  for (auto& p : c->md_peers) {
    p.send_idx.release(); p.sendbuf.release(); p.recvbuf.release();
    p.add.seg_off.release(); p.add.seg_dst.release(); p.add.idx.release();
  }
  c->md_peers.clear();
  c->md_peers.resize(n_peers);
  size_t so = 0, ro = 0;
  for (int i = 0; i < n_peers; i++) {
    pp_context::MdPeer& p = c->md_peers[i];
    PP_REQUIRE(peer_rank[i] >= 0 && peer_rank[i] < c->comm.nranks && peer_rank[i] != c->comm.rank, "pp_md_exchange_setup: bad peer rank %d", peer_rank[i]);
    p.rank = peer_rank[i];
    p.n_send = send_count[i];
    p.n_recv = recv_count[i];
    for (int k = 0; k < p.n_send; k++)
      PP_REQUIRE(send_idx_flat[so + k] >= nlocal && send_idx_flat[so + k] < ntotal, "pp_md_exchange_setup: send index %d is not a ghost", send_idx_flat[so + k]);
    if (p.send_idx.ensure(std::max(p.n_send, 1), 0, c->stream) || p.sendbuf.ensure(3 * (size_t)std::max(p.n_send, 1), 0, c->stream) ||
        p.recvbuf.ensure(3 * (size_t)std::max(p.n_recv, 1), 0, c->stream))
      return 1;
    if (p.n_send) PP_CUDA(cudaMemcpy(p.send_idx.p, send_idx_flat + so, p.n_send * sizeof(int), cudaMemcpyHostToDevice));
    if (build_seglist(c, p.add, p.n_recv, nullptr, recv_dst_flat + ro, nlocal, "received")) return 1;
    so += p.n_send;
    ro += p.n_recv;
  }
  c->md_exchange_set = true;
  return 0;
}

// forces of ghosts back to their owners (f_dev[ntotal][3], accumulated in place), energy summed over the ranks
extern "C" int pp_md_exchange_device(pp_context* c, double* f_dev, double* etotal_dev)
{
  PP_REQUIRE(c && c->md_exchange_set, "pp_md_exchange_device: call pp_md_exchange_setup first");
  PP_REQUIRE(f_dev, "pp_md_exchange_device: NULL force array");
  PP_CUDA(cudaSetDevice(c->device));
  cudaStream_t st = c->stream;
  PP_CUDA(cudaEventRecord(c->ev[4], st));
  const int np = (int)c->md_peers.size();
  std::vector<int> ranks(np);
  std::vector<double*> sb(np), rb(np);
  std::vector<size_t> sn(np), rn(np);
  for (int i = 0; i < np; i++) {
    pp_context::MdPeer& p = c->md_peers[i];
    if (ppk_pack_forces(f_dev, p.send_idx.p, p.n_send, p.sendbuf.p, st, &c->launches)) return 1;
    ranks[i] = p.rank; sb[i] = p.sendbuf.p; rb[i] = p.recvbuf.p; sn[i] = 3 * (size_t)p.n_send; rn[i] = 3 * (size_t)p.n_recv;
  }
  if (c->comm.sendrecv(np, ranks.data(), sb.data(), sn.data(), rb.data(), rn.data(), st)) return 1;
  // own ghosts first, then the peers in the order given: a fixed order, so the forces are reproducible
  if (ppk_segment_add(f_dev, c->md_self.seg_off.p, c->md_self.seg_dst.p, c->md_self.n_seg, f_dev, c->md_self.idx.p, st, &c->launches)) return 1;
  for (int i = 0; i < np; i++) {
    pp_context::MdPeer& p = c->md_peers[i];
    if (ppk_segment_add(f_dev, p.add.seg_off.p, p.add.seg_dst.p, p.add.n_seg, p.recvbuf.p, p.add.idx.p, st, &c->launches)) return 1;
  }
  if (etotal_dev && c->comm.nranks > 1) {
    // each rank's total carries the output shift (pair_nn.cpp:175-179 adds it on rank 0 only): keep it once
    if (c->comm.rank != 0 && c->out_mean != 0.0 && ppk_add_scalar(etotal_dev, -c->out_mean, st, &c->launches)) return 1;
    if (c->comm.allreduce_device(etotal_dev, 1, st)) return 1;
  }
  PP_CUDA(cudaEventRecord(c->ev[5], st));
  return 0;
}

// One MD step of a partitioned box through host buffers: x[ntotal][3] uploaded, forces computed on locals and ghosts,
// ghost forces returned to their owners (on this rank and on the peers), the forces of the nlocal LOCAL atoms
// downloaded (overwritten, not added), energies downloaded; *etotal_out is the sum over all ranks.
extern "C" int pp_md_step_exchange(pp_context* c, const double* x, double* f_local_out, double* eatom_out, double* etotal_out)
{
  PP_REQUIRE(c && c->md_list_resident && c->md_exchange_set, "pp_md_step_exchange: needs pp_md_build_neighbors and pp_md_exchange_setup");
  PP_REQUIRE(x, "pp_md_step_exchange: NULL positions");
  PP_CUDA(cudaSetDevice(c->device));
  cudaStream_t st = c->stream;
  const int ntotal = c->md_ntotal, nlocal = c->md_nlocal;
  if (c->d_etotal.ensure(1, 0, st)) return 1;
  PP_CUDA(cudaMemcpyAsync(c->d_x.p, x, 3 * (size_t)ntotal * sizeof(double), cudaMemcpyHostToDevice, st));
  PP_CUDA(cudaMemsetAsync(c->d_f.p, 0, 3 * (size_t)ntotal * sizeof(double), st));
  if (md_run(c, c->d_x.p, c->d_f.p)) return 1;
  if (md_energy_total(c, c->d_etotal.p)) return 1;
  if (pp_md_exchange_device(c, c->d_f.p, c->d_etotal.p)) return 1;
  if (f_local_out) PP_CUDA(cudaMemcpyAsync(f_local_out, c->d_f.p, 3 * (size_t)nlocal * sizeof(double), cudaMemcpyDeviceToHost, st));
  if (eatom_out) PP_CUDA(cudaMemcpyAsync(eatom_out, c->d_eatom.p, nlocal * sizeof(double), cudaMemcpyDeviceToHost, st));
  double et = 0.0;
  PP_CUDA(cudaMemcpyAsync(&et, c->d_etotal.p, sizeof(double), cudaMemcpyDeviceToHost, st));
  PP_CUDA(cudaStreamSynchronize(st));
  tock(c, 0, 1, 4); tock(c, 1, 2, 5); tock(c, 2, 3, 6); tock(c, 4, 5, 7);
  if (etotal_out) *etotal_out = et;
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
extern "C" int pp_comm_unique_id(void* id) { return PPComm::unique_id(id); }
extern "C" int pp_comm_init(pp_context* c, int nranks, int rank, const void* id)
{
  PP_REQUIRE(c, "null context");
  PP_CUDA(cudaSetDevice(c->device));
  return c->comm.init(nranks, rank, id);
}
extern "C" int pp_comm_allreduce_host(pp_context* c, double* buf, int n)
{
  PP_REQUIRE(c && (buf || n == 0) && n >= 0, "pp_comm_allreduce_host: bad arguments");
  PP_CUDA(cudaSetDevice(c->device));
  return c->comm.allreduce_host(buf, n, c->stream);
}
extern "C" const char* pp_comm_path(pp_context* c) { return c ? c->comm.path() : ""; }
extern "C" int pp_comm_rank(pp_context* c) { return c ? c->comm.rank : 0; }
extern "C" int pp_comm_size(pp_context* c) { return c ? c->comm.nranks : 1; }
