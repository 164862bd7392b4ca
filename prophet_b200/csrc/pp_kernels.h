// Internal kernel-launcher interface between pp_api.cu and the kernel translation units.
#pragma once
#include "pp_common.cuh"

// ---- descriptors / forces -------------------------------------------------------------------------------
struct PeriodicArgs {
  const DescSpec* spec;            // device
  long long n_atoms;               // atoms handled by this launch
  long long atom_begin;            // first (global) atom of this launch
  const int* atom_struct;          // [total atoms] structure index
  const long long* struct_atom0;   // [n_struct+1]
  const double* FtoA;              // [n_struct][9]   Structure::FtoA = cell^T (Structure.cpp:1186)
  const int* mxyz;                 // [n_struct][3]   image search box (Structure.cpp:1300-1302)
  const double* frac;              // [total atoms][3]
  const int* elem;                 // [total atoms]
  double* G;                       // [total atoms][NG]
  int cap;
  int* overflow;
};

struct MdArgs {
  const DescSpec* spec;
  int nlocal;
  const double* x;                 // [ntotal][3]
  const int* elem;                 // [ntotal] element index
  const long long* neigh_offsets;  // [nlocal+1]
  const int* neigh;                // flat full list
  double* G;                       // [nlocal][NG]     (desc)
  const double* dEdG;              // [nlocal][NG]     (force)
  double* f;                       // [ntotal][3]      (force, accumulated)
  int cap;
  int* overflow;
};

int ppk_desc_periodic(const DescSpec& hs, PeriodicArgs a, int* cap_io, cudaStream_t st, long long* launches);
int ppk_neighbors_periodic(const DescSpec& hs, PeriodicArgs a, int s, int natom, int pass, int* counts, const int* offsets,
                           int* out_j, int* out_t, int cap, cudaStream_t st, long long* launches);
int ppk_max_cap(const DescSpec& hs, bool force);
int ppk_md_desc(const DescSpec& hs, MdArgs a, int* cap_io, cudaStream_t st, long long* launches);
int ppk_md_force(const DescSpec& hs, MdArgs a, int cap, cudaStream_t st, long long* launches);

// ---- per-element MLP --------------------------------------------------------------------------------------
struct NetDev {
  int n_layers;                       // weight layers (hidden + output)
  int sizes[PP_MAX_LAYERS + 1];       // sizes[0] = NG ... sizes[n_layers] = 1
  int w_off[PP_MAX_LAYERS];           // offset of layer l's [n_out][n_in+1] block in the flat parameter vector
  int node_off[PP_MAX_LAYERS];        // offset of layer l's nodes in the transfer-id list
  int P;
  unsigned tanh_layers;               // bit l set: every node of layer l (1-based, hidden) has the "tanh" transfer (id 1)
  unsigned spline_layers;             // bit l set: every node of layer l has the "tanh_spline" transfer (id 2)
  const double* params;               // device, this element's parameters
  const int* tf;                      // device, transfer id per node
  const double* mean;                 // device [NG]
  const double* scale;                // device [NG]
};

enum { PP_MLP_FORWARD = 0, PP_MLP_TRAIN = 1, PP_MLP_INPUT_GRAD = 2, PP_MLP_FUSED = 3,
       PP_MLP_TRAIN_STORED = 4 };      // template mode of mlp_fused2_kernel only: TRAIN reading the activation store

// fused MD step (md_fused_kernel): the per-element networks of the atom-centred pass
struct MdNets {
  NetDev net[PP_MAX_ELEM];
  const double* WT[PP_MAX_ELEM];   // per element: every weight layer transposed, [n_in + 1][n_out] at the layer's w_off
  double out_scale;
  const double* tanh_table;
  double* eatom;                   // [nlocal] out
  double* dEdG_store;              // [nlocal][NG] out, or nullptr (parity hook)
  double* G_store;                 // [nlocal][NG] out, or nullptr (parity hook)
  int cap_lo;                      // only atoms with cap_lo < neighbours <= cap are processed (re-run after an overflow)
};
bool ppk_md_fused_ok(const NetDev& net);
int ppk_md_fused(const DescSpec& hs, MdArgs a, const MdNets& nets, int* cap_io, cudaStream_t st, long long* launches);


struct MlpArgs {
  NetDev net;
  const double* G;                    // [atoms][NG]
  const int* rows;                    // [M] atom indices of this element (nullptr = identity)
  int prenorm;                        // 1: G holds (x - mean) / scale already (pp_api.cu: normalised copy), the kernels skip NN.cpp:647-651
  int M;
  const int* atom_struct;             // [atoms]                    (train)
  const double* cstruct;              // [n_struct] 2*Err*out_scale (train)
  double out_scale;
  double* E_atom;                     // [atoms]  out (forward)
  double* dEdG;                       // [atoms][NG] out (input grad)
  double* partial;                    // [n_cta][P_total+3] out (train); this element's slice starts at p_off;
                                      //   fused mode also fills the 3 tail columns (sse, vsse, nval)
  int P_total, p_off;
  const double* tanh_table;           // [1002 + 4000], see pp_fastmath.cuh
  const void* dw_tiles;               // device table of weight-gradient tiles (ppk_mlp_num_dw_tiles)
  const void* dw_tiles2;              // device work table of the two-CTA fused kernel (ppk_mlp2_table), or nullptr
  double* scratch2;                   // two-CTA form: derivative scratch, ppk_mlp2_scratch_doubles(net) doubles
  // two-CTA form, two-phase epoch with stored activations (several elements): hidden activations and exact derivatives
  // [M][store_w], layer l's n_l columns after those of the layers before it.  FORWARD writes them, TRAIN reads them
  // instead of running the forward pass again; nullptr = recompute
  double* hstore;
  double* tstore;
  int store_w;
  // fused single-pass epoch (single-element systems, tiles made of whole structures)
  int n_tiles;
  const int* tile_struct0;            // [n_tiles+1] first structure of each tile
  const long long* struct_atom0;      // [n_struct+1]
  const double* target;               // [n_struct]
  const int* tag;                     // [n_struct]
  double out_mean;
  double* E_struct;                   // [n_struct] out
};

// returns the number of CTAs the train launch will use (rows of `partial`)
int ppk_mlp_max_ctas(void);
int ppk_mlp_single_ctas(void);
// atoms per tile the kernel will use for this net (0 = does not fit), and its weight-gradient tile table
int ppk_mlp_tile_atoms(const NetDev& net);
int ppk_mlp_num_dw_tiles(const NetDev& net, void* host_tiles_out);
int ppk_mlp_dw_tile_bytes(void);
// two-CTA fused form: work table (0 = network not eligible) and launch; *n_cta = partial rows written
int ppk_mlp2_table(const NetDev& net, void* host_out, int* n_entries);
size_t ppk_mlp2_scratch_doubles(const NetDev& net);
int ppk_mlp_launch_fused(const MlpArgs& a, cudaStream_t st, long long* launches, int* n_cta);
bool ppk_mlp_two_cta_ok(const NetDev& net);
bool ppk_mlp2_available(void);          // the driver entry point the two-CTA launches need was found
int ppk_mlp_launch2(int mode, bool allow_two_cta, const MlpArgs& a, cudaStream_t st, long long* launches, int* n_cta);
int ppk_mlp_launch(int mode, const MlpArgs& a, cudaStream_t st, long long* launches);

// tensor map of a row-major [rows][cols] FP64 array with box (box_cols x box_rows); out_map is a CUtensorMap (cached)
bool ppk_tiled_map(const double* base, int cols, long long rows, int box_cols, int box_rows, void* out_map);
int ppk_sm_count(void);

// ---- two-phase epoch with stored activations (pp_mlpw.cu): systems of several elements, wide / deep networks ----
// H / T: activations and exact transfer-function derivatives of the hidden layers for the atoms of ONE element, layer
// after layer: layer l (1 .. n_layers-1) is a row-major [M][sizes[l]] block at offset M * (sizes[1] + .. + sizes[l-1]),
// rows in the order of MlpArgs::rows.  H == nullptr: nothing is stored (forward pass only).
struct MlpwStore {
  double* H;
  double* T;
};
bool ppk_mlpw_ok(const NetDev& net);                                   // network shape eligible for these kernels
int ppk_mlpw_table(const NetDev& net, void* host_out, int* n_entries); // weight-gradient work table (MlpArgs::dw_tiles2)
int ppk_mlpw_store_width(const NetDev& net);                           // doubles of H (and of T') per atom
int ppk_mlpw_forward(const MlpArgs& a, const MlpwStore& st, cudaStream_t stream, long long* launches);
int ppk_mlpw_backward(const MlpArgs& a, const MlpwStore& st, cudaStream_t stream, long long* launches, int* n_cta);

// ---- FP32 mode (pp_mlp32.cu): descriptors kept as FP32, single-pass training epoch on 3xTF32 tensor-core products ----
bool ppk_tiled_map_f32(const float* base, int cols, int pitch_floats, long long rows, int box_cols, int box_rows, void* out_map);
bool ppk_mlp32_ok(const NetDev& net);
int ppk_mlp32_pitch(const NetDev& net);                 // row pitch (floats) of the FP32 descriptor array
int ppk_mlp32_tile_atoms(void);
int ppk_mlp32_table(const NetDev& net, void* host_out, int* n_entries);      // returns the entry size in bytes (0: not eligible)
// FP64 descriptors -> their normalised FP64 copy (x - mean[e]) * inv_scale[e], the arithmetic the kernels apply per tile
int ppk_g_normalise(const double* G, int NG, long long n_atoms, const int* elem, const double* mean, const double* inv_scale, double* out,
                     cudaStream_t st, long long* launches);
int ppk_g_to_f32(const double* G, int NG, int NGp, long long n_atoms, const int* elem, const double* mean, const double* scale, float* out,
                 cudaStream_t st, long long* launches);
int ppk_mlp32_launch_fused(const MlpArgs& a, const float* G32, long long n_atoms, cudaStream_t st, long long* launches, int* n_cta);
// the same epoch with the forward and delta products on tcgen05.mma (pp_mlp32u.cu): eligibility (PP_FP32_NO_UMMA=1 switches
// it off), tile size, launch (same work table as ppk_mlp32_table), and the flag a launch raises when a tensor-pipe batch
// never completed (reading it synchronises the stream)
bool ppk_mlp32u_ok(const NetDev& net);
int ppk_mlp32u_tile_atoms(void);
int ppk_mlp32u_launch_fused(const MlpArgs& a, const float* G32, long long n_atoms, cudaStream_t st, long long* launches, int* n_cta);
int ppk_mlp32u_failed(cudaStream_t st);

// per-structure energies / errors (Potential.cpp:641-668)
struct StructArgs {
  int n_struct;
  const long long* struct_atom0;
  const double* E_atom;
  const double* target;
  const int* tag;
  double out_mean, out_scale;
  double* E_struct;     // [n_struct]
  double* cstruct;      // [n_struct]
  double* result_tail;  // [3] sse, vsse, nval
  double* scratch;      // [3 * ppk_struct_errors_blocks(n_struct)] per-block loss terms
  unsigned* ticket;     // zero before the first launch; the kernel leaves it at zero
};
int ppk_struct_errors_blocks(int n_struct);
int ppk_struct_errors(const StructArgs& a, cudaStream_t st, long long* launches);
int ppk_reduce_partials(const double* partial, int n_cta, int stride, int ncols, double* out, cudaStream_t st, long long* launches);

// preconditioning column statistics (Structure::mean / variance, Structure.cpp:1443-1476)
// pass 0: sums[k] = sum G[a][k]; pass 1: sums[k] = sum (G[a][k]-mean[k])^2 ; over atoms in rows[] whose structure is 'train'
int ppk_column_stats(const double* G, int NG, const int* rows, int M, const int* atom_struct, const int* tag,
                     const double* mean, int pass, double* partial, int n_cta, double* sums, long long* count,
                     cudaStream_t st, long long* launches);

void ppk_make_tanh_table(std::vector<double>& table);

// ---- MD harness: GPU cell list and full neighbour list (pp_cell.cu) ------------------------------------------
struct CellGrid {
  double lo[3], inv[3];   // cell index = (x - lo) * inv, clamped
  int n[3];
};
struct CellWork {
  int* cell_of;           // [ntotal]
  int* cell_count;        // [ncells]
  long long* cell_start;  // [ncells+1]
  int* sorted;            // [ntotal] atoms grouped by cell, ascending index inside a cell
  int* counts;            // [nlocal] neighbours per local atom
  long long* block_sums;  // scan scratch
  long long* total;       // [1]
};
int ppk_build_cell_neighbors(const double* x_dev, int nlocal, int ntotal, const CellGrid& g, double cutneigh, CellWork& w,
                             cudaStream_t st, long long* launches);
int ppk_scan_counts(const int* counts, long long n, long long* offsets, CellWork& w, cudaStream_t st, long long* launches);
int ppk_fill_cell_neighbors(const double* x_dev, int nlocal, const CellGrid& g, double cutneigh, CellWork& w, const long long* offsets,
                            int* neigh_out, cudaStream_t st, long long* launches);
int ppk_sum(const double* v, long long n, double add, double* partial, double* out, cudaStream_t st, long long* launches);

// ---- reverse communication of ghost forces (pp_exchange.cu) -----------------------------------------------------
int ppk_pack_forces(const double* f, const int* idx, int n, double* out, cudaStream_t st, long long* launches);
int ppk_add_scalar(double* v, double a, cudaStream_t st, long long* launches);
int ppk_segment_add(double* f, const int* seg_off, const int* seg_dst, int n_seg, const double* base, const int* idx, cudaStream_t st,
                    long long* launches);
