/*
 * prophet_b200.h -- C ABI of the B200 device layer for PROPhet's Behler-Parrinello hot path.
 *
 * The reference (biklooost/PROPhet) has no C ABI of its own: the seams this layer sits under are the C++
 * classes Structure / Neural_network / Potential and the LAMMPS pair style PairNN.  Each entry point below
 * names the reference function(s) it replaces (file:line relative to the reference's src/).  The host-side
 * C++ classes of this repository (prophet_b200/host: Setup, Functional_params, Potential, Optimizer, Parallel,
 * PairNN) call ONLY these functions; so can any other host language through its FFI.
 *
 * Conventions
 *   - every function returns 0 on success, non-zero on error; pp_last_error() gives the message
 *     (the reference's convention is "any error is fatal", Error.cpp:49-81 -- callers may simply abort);
 *   - all pointers are HOST pointers unless the name ends in _dev; buffers are caller-owned;
 *   - one pp_context per (process, GPU); a context is not thread-safe (the reference is single-threaded per rank);
 *   - all arithmetic is IEEE double (the reference's REAL, Common.h:59); indices are int unless stated;
 *   - elements are referred to by their INDEX in the ascending-atomic-number list given to pp_set_descriptors
 *     (Potential::get_all_atom_types, Potential.cpp:888-917; Structure::init_type_G_map, Structure.cpp:1491-1512);
 *   - there is no CPU fallback: without a CUDA device pp_create fails.
 */
#ifndef PROPHET_B200_H
#define PROPHET_B200_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct pp_context pp_context;

/* transfer-function ids (Neural_network_node::set_transfer_function, NN_node.cpp:120-171; "random" is not
 * supported: it is not a function) */
enum pp_transfer {
  PP_TF_LINEAR = 0, PP_TF_TANH = 1, PP_TF_TANH_SPLINE = 2, PP_TF_TANH_CRAZY = 3, PP_TF_CUBIC = 4,
  PP_TF_PULSE = 5, PP_TF_SIN = 6, PP_TF_PERIODIC_TANH = 7, PP_TF_RELU = 8
};

/* structure tags (System::train, System.cpp:59,117-119: "train" | "val" | anything else) */
enum pp_tag { PP_TAG_TRAIN = 0, PP_TAG_VAL = 1, PP_TAG_OTHER = 2 };

const char* pp_last_error(void);
int pp_version(void);

/* ---- context --------------------------------------------------------------------------------------------- */
/* Create a context on CUDA device `device`.  Replaces nothing in the reference (it has no device); plays the role of
 * the Potential constructor's resource set-up.
 * fp_mode 64: every kernel in IEEE double, the reference's REAL (Common.h:59); results within relative 1e-10.
 * fp_mode 32: the reference's USE_FLOAT switch (Common.h:53-61; its own float build does not compile, so the FP64
 *   reference at relative 1e-5 is the yardstick).  Training epochs of single-element systems with tanh hidden layers
 *   (one or two, up to 64 wide, structures of up to 256 atoms) run in single precision: the descriptors are kept as
 *   normalised FP32 and the layer products are 3xTF32 tensor-core products accumulated in FP32 (forward and delta
 *   products on tcgen05.mma with TMEM accumulators when every structure has at most 128 atoms and the layers are at most
 *   63 wide, mma.sync otherwise; PP_FP32_NO_UMMA=1 keeps mma.sync); structure energies,
 *   errors, loss terms and the final gradient sum stay FP64 (pp_epoch_form == 4).  Descriptors, neighbour lists,
 *   evaluation, MD and every other system keep the FP64 kernels in this mode. */
int pp_create(int device, int fp_mode, pp_context** out);
void pp_destroy(pp_context* ctx);

/* ---- descriptors: Structure::init_G / Allocate_G (Structure.cpp:779-960) ----------------------------------- */
/* Rows: G1 [Rc, fc], G2 [Rc, fc, eta, Rs], G3/G4 [Rc, fc, eta, xi, lambda] (Structure.cpp:786-791), row-major.
 * n_elem elements in ascending Z.  NG = n_elem*(nG1+nG2) + n_elem*(n_elem+1)/2*(nG3+nG4) (Structure.cpp:859-860). */
int pp_set_descriptors(pp_context* ctx, int n_elem, int nG1, const double* G1, int nG2, const double* G2,
                       int nG3, const double* G3, int nG4, const double* G4, double Rcut);
int pp_num_descriptors(pp_context* ctx);

/* ---- networks: Neural_network ctor/load/set_state (NN.cpp:51-131,158-200,805-819), Potential::insert_atom_type
 *      (Potential.cpp:1214-1253), Network::set_means/set_variances (NN.h:93-100) ------------------------------ */
/* sizes[n_layers+1] = {NG, hidden..., 1}; node_transfer: one id per node over all layers (last node linear);
 * params: layer-major, node-major, [w_0..w_{n-1}, b] per node (NN.cpp:83-94) -- the same flat vector
 * Optimizer::get_parameters (Optimizer.cpp:329-348) produces; in_mean/in_scale: NG each. */
int pp_set_network(pp_context* ctx, int elem, int n_layers, const int* sizes, const int* node_transfer,
                   const double* params, const double* in_mean, const double* in_scale);
/* output preconditioning: Potential::output_mean / output_variance (Potential.cpp:461-500) */
int pp_set_output_scaling(pp_context* ctx, double out_mean, double out_scale);
/* total parameter count over elements (ascending element index = Potential::indices order, Potential.cpp:190-194) */
int pp_num_parameters(pp_context* ctx);
/* Optimizer::set_parameters (Optimizer.cpp:303-317): all elements' parameters, concatenated */
int pp_set_parameters(pp_context* ctx, const double* params);
int pp_get_parameters(pp_context* ctx, double* params);
int pp_set_input_scaling(pp_context* ctx, int elem, const double* in_mean, const double* in_scale);

/* ---- training set: System/Structure construction + Potential ctor loop (Potential.cpp:113-150) -------------- */
/* Cartesian -> fractional + single wrap: Structure::Set_Convert + Normalize (Structure.cpp:1159-1279).
 * cell = rows a,b,c.  Bit-identical to the reference (host arithmetic, no FMA contraction). */
int pp_cart_to_frac(const double* cell, const double* cart, int n, double* frac_out);
/* The same for n_struct structures in one call (the loop over systems of the Potential constructor,
 * Potential.cpp:113-150): natoms[n_struct], cells[n_struct][9], cart / frac_out[sum natoms][3]. */
int pp_cart_to_frac_batch(int n_struct, const int* natoms, const double* cells, const double* cart, double* frac_out);
/* Append n_struct periodic structures and compute their descriptors on the GPU
 * (Set_Neighbor :1287-1353 + Calc_G periodic branch :985-1071, fused; never materialised on the host).
 *   natoms[n_struct]; cells[n_struct][9] rows a,b,c; frac[sum natoms][3] fractional coordinates as
 *   Structure::pos holds them after Normalize; elem[sum natoms] element index; targets[n_struct]; tags[n_struct]. */
int pp_add_structures(pp_context* ctx, int n_struct, const int* natoms, const double* cells, const double* frac,
                      const int* elem, const double* targets, const int* tags);
int pp_clear_structures(pp_context* ctx);
/* Structure::init_G run again on every resident structure after pp_set_descriptors changed the row PARAMETERS
 * (same counts, same Rcut): the per-step descriptor refresh of Potential::optimize_Gs (Potential.cpp:806-808). */
int pp_recompute_descriptors(pp_context* ctx);
int pp_num_structures(pp_context* ctx);
long long pp_num_atoms(pp_context* ctx);
/* parity hooks.  G_out[NG][natoms(s)] in the reference's Structure::G layout. */
int pp_get_descriptors(pp_context* ctx, int s, double* G_out);
/* Neighbour list of structure s exactly as Structure::Set_Neighbor builds it: neigh_offset[natoms+1],
 * neigh_j (unique j ascending), image_offset[#pairs+1], image_t (translation-table indices ascending).
 * Two-call protocol: pass NULL arrays to obtain the counts in n_pairs / n_images. */
int pp_get_neighbors(pp_context* ctx, int s, int* neigh_offset, int* neigh_j, int* image_offset, int* image_t,
                     int* n_pairs, int* n_images);

/* ---- Potential::Precondition input part (Potential.cpp:404-455; Structure::mean/variance :1443-1476) ------- */
/* Computes per-element mean and sqrt-variance (floored at 1e-5) over atoms of 'train' structures and installs
 * them in the networks.  With a communicator set (pp_comm_init) sums and counts are all-reduced first. */
int pp_precondition(pp_context* ctx, int input_precondition);
int pp_get_input_scaling(pp_context* ctx, int elem, double* in_mean, double* in_scale);

/* ---- one training epoch: the body of Potential::train's loop (Potential.cpp:626-670) ----------------------- */
/* params_in (may be NULL = keep current) -> forward all atoms, per-structure error, backprop to all parameters.
 * grad_out[P] = sum over train structures of 2*Err*dE/dtheta; *sse, *vsse, *nval as the reference accumulates them.
 * With a communicator the four outputs are summed over ranks (replaces Optimizer.cpp:887-888,969-970). */
int pp_epoch(pp_context* ctx, const double* params_in, double* grad_out, double* sse, double* vsse, long long* nval);
/* the same pass with parameters already on the device and results left on the device (no host copies):
 * result layout on device = [grad(P), sse, vsse, nval] doubles. */
int pp_epoch_device(pp_context* ctx);
int pp_epoch_result_dev(pp_context* ctx, double** result_dev, int* n);
/* Diagnostic: which kernel form the last epoch ran.  0 = two launches per element (forward, then backward with the
 * forward pass recomputed: any system), 1 = single-pass fused kernel, one CTA per SM, 2 = single-pass fused kernel with
 * in-place deltas, two CTAs per SM (single element, tanh / tanh_spline hidden nodes), 3 = two launches per element with
 * the activations kept in HBM between them (several elements, or networks too large for the single-pass kernels; even
 * layer widths up to 64, at most four hidden layers), 4 = the single-precision single-pass kernel of fp_mode 32,
 * 5 = form 0 through the two-CTA kernels with the activations kept in HBM between the two launches instead of the
 * forward pass being run again (several elements, networks of one or two hidden layers up to 64 wide).
 * Same arithmetic in forms 0 - 3 and 5, different summation order.
 * -1 before the first epoch.  Environment: PP_MLP_ONE_CTA=1 keeps form 2 from being chosen, PP_MLP_NO_STORE=1 forms 3 and 5;
 * PP_MLP_NO_PRENORM=1 makes the kernels normalise the descriptors per tile instead of reading the library's normalised
 * copy (same results bit for bit; the copy costs NG doubles per atom of device memory and is dropped by itself when that
 * allocation fails).
 * No reference counterpart. */
int pp_epoch_form(pp_context* ctx);
/* Potential::evaluate (Potential.cpp:515-551): E_out[n_struct] = sum_atoms E_i + out_mean (every tag) */
int pp_evaluate(pp_context* ctx, double* E_out);
/* per-atom energies of structure s (Neural_network::evaluate_MD(int), NN.cpp:585-612) */
int pp_get_atom_energies(pp_context* ctx, int s, double* E_atom_out);

/* ---- LAMMPS pair_style nn: PairNN::compute (pair_nn.cpp:112-188) -------------------------------------------- */
/* type_to_elem[ntypes+1]: LAMMPS type id -> element index (Potential::insert_atom_type, Potential.cpp:1236-1250) */
int pp_md_set_types(pp_context* ctx, int ntypes, const int* type_to_elem);
/* x[ntotal][3], type[ntotal], full neighbour list of the nlocal local atoms (ilist must be the identity,
 * pair_nn.cpp:151-157): numneigh[nlocal], neigh_offsets[nlocal+1] (64-bit), neigh_flat.
 * f_inout[ntotal][3] is ADDED to (locals and ghosts, Structure::Get_Forces Structure.cpp:429-713);
 * eatom_out[nlocal] (may be NULL) receives per-atom energies; *etotal_out = sum + out_mean (pair_nn.cpp:156-179).
 * G_out / dEdG_out [nlocal][NG] (may be NULL) are parity hooks for Calc_G (:1073-1145) and
 * Neural_network::evaluate_MD(dE_dG) (NN.cpp:516-570). */
int pp_md_compute(pp_context* ctx, const double* x, const int* type, int nlocal, int nghost,
                  const long long* neigh_offsets, const int* neigh_flat,
                  double* f_inout, double* eatom_out, double* etotal_out, double* G_out, double* dEdG_out);
/* Device-resident MD for harnesses that keep the list on the GPU (the C5 benchmark):
 * build a full neighbour list with a GPU cell list using LAMMPS' test rsq < cutneigh^2 over locals+ghosts. */
int pp_md_build_neighbors(pp_context* ctx, const double* x, const int* type, int nlocal, int nghost, double cutneigh);
int pp_md_get_neighbors(pp_context* ctx, long long* neigh_offsets, int* neigh_flat, long long* n_total);
/* one step on the resident list: uploads x, downloads f (added into f_inout) and energies */
int pp_md_step(pp_context* ctx, const double* x, double* f_inout, double* eatom_out, double* etotal_out);
/* the same with x / f already on the device (x_dev[ntotal][3], f_dev[ntotal][3] accumulated into) */
int pp_md_step_device(pp_context* ctx, const double* x_dev, double* f_dev, double* etotal_dev);

/* ---- partitioned box: the host engine's reverse communication ------------------------------------------------------ */
/* PairNN::compute leaves forces on ghost atoms (Structure::Get_Forces writes f[j] for every neighbour, pair_nn.cpp:182)
 * and counts on LAMMPS' Comm::reverse_comm to return them to the atoms that own them.  When a caller drives a box cut
 * over the ranks of a communicator itself (pp_comm_init), these calls play that part on the GPU: ghost forces are packed
 * by a kernel, travel by NCCL send/recv on the context's stream and are added in a fixed order (reproducible forces).
 *   own ghosts : n_self pairs (ghost index self_src[k] >= nlocal, local index self_dst[k]): periodic images of this
 *                rank's own atoms;
 *   peers      : for peer i (rank peer_rank[i]) send_count[i] ghost indices in send_idx_flat (concatenated over peers),
 *                sent in that order, and recv_count[i] local indices in recv_dst_flat: where the values that peer sends
 *                (in ITS order) are added.  The lists are whatever the caller's decomposition dictates.
 * Call after pp_md_build_neighbors (which fixes nlocal / nghost). */
int pp_md_exchange_setup(pp_context* ctx, int n_self, const int* self_src, const int* self_dst, int n_peers, const int* peer_rank,
                         const int* send_count, const int* send_idx_flat, const int* recv_count, const int* recv_dst_flat);
/* f_dev[ntotal][3] (device): ghost forces added onto their owners here and on the peers; etotal_dev (device, may be NULL):
 * this rank's total as pp_md_step_device leaves it -> the sum over ranks (the output shift counted once, as
 * pair_nn.cpp:175-179 adds it on rank 0 only). */
int pp_md_exchange_device(pp_context* ctx, double* f_dev, double* etotal_dev);
/* pp_md_step for a partitioned box: x[ntotal][3] uploaded, forces on locals and ghosts, exchange, then the forces of
 * the nlocal LOCAL atoms written to f_local_out[nlocal][3] (overwritten, LAMMPS having cleared them), per-atom energies
 * and the total over all ranks. */
int pp_md_step_exchange(pp_context* ctx, const double* x, double* f_local_out, double* eatom_out, double* etotal_out);

/* ---- Parallel (Parallel.h:93-214): NCCL replaces MPI_Reduce/MPI_Bcast of Optimizer::update_network ---------- */
/* unique_id: 128-byte ncclUniqueId made by pp_comm_unique_id on rank 0 and distributed by the launcher */
int pp_comm_unique_id(void* unique_id_128);
int pp_comm_init(pp_context* ctx, int nranks, int rank, const void* unique_id_128);
int pp_comm_rank(pp_context* ctx);
/* Diagnostic: how the per-epoch sum over ranks travels -- "peer memory" (every rank's reduction kernel stores its sums
 * into receive slots in each peer's memory over NVLink and every rank adds the slots in rank order: one-shot, latency
 * bound), "nccl" (ncclAllReduce; also chosen by PP_COMM_NCCL_ONLY=1, when peers cannot be mapped, or beyond 8 ranks) or
 * "single rank".  Decided at the first epoch.  No reference counterpart. */
const char* pp_comm_path(pp_context* ctx);
int pp_comm_size(pp_context* ctx);
/* in-place sum over ranks of n host doubles, result on every rank: Parallel::Reduce(..., MPI_SUM) followed by
 * Parallel::Bcast (Parallel.h:93-150), the pattern of every small collective in Potential / Setup; identity for one rank */
int pp_comm_allreduce_host(pp_context* ctx, double* buf, int n);

/* ---- descriptor statistics without allocation ------------------------------------------------------------------ */
/* No counterpart in the reference (it has no device memory to manage).
 * For callers that must not allocate device memory -- a host program whose own pool already owns the GPU's memory,
 * or a time-critical loop that cannot tolerate the device-wide synchronisation cudaMalloc/cudaFree may cause.
 * Computes, for a row-major device matrix [n_rows][n_cols] (e.g. descriptors, one atom per row):
 *     squared == 0 :  sums_out[k] = sum_r M[r][k]
 *     squared != 0 :  sums_out[k] = sum_r (M[r][k] - shift[k])^2      (shift_dev may be NULL = 0)
 * i.e. the two reductions of Potential::Precondition (Potential.cpp:404-455), with the intermediate per-block sums
 * in a fixed 256 KiB array inside the library instead of allocated scratch.  Limits and rules:
 *   - n_cols <= 512, otherwise the call fails with an error (use pp_precondition, which allocates);
 *   - the array is held by every GPU that loads the library until the process exits;
 *   - one array per GPU, shared by all calls: calls that could overlap on the same GPU must be serialised by the caller;
 *   - the first call on a GPU resolves the array's address (a set-up step): make it before any stream capture, and
 *     do not let two first calls for the same GPU run at the same time on different threads;
 *   - the caller's current device is the same after the call as before it.
 * All pointers are DEVICE pointers; `stream` is a cudaStream_t (NULL = default stream); the call does not synchronise. */
int pp_column_sums_static(int device, const double* matrix_dev, long long n_rows, int n_cols,
                          const double* shift_dev, int squared, double* sums_out_dev, void* stream);

/* ---- instrumentation --------------------------------------------------------------------------------------- */
/* number of kernels this library launched since the context was created */
long long pp_kernel_launches(pp_context* ctx);
/* device-side time (ms, CUDA events on the context's stream) of the last call's phases:
 * which: 0 = descriptors (pp_add_structures), 1 = epoch forward, 2 = epoch backward, 3 = reduction of the per-CTA
 *        partial gradients + sum over ranks,
 *        4 = md descriptors, 5 = md network, 6 = md forces (with PP_MD_FUSED=1 the step runs as ONE kernel per atom
 *        shell -- slower than the three kernels on B200, kept as a measured alternative -- and phase 6 is that kernel),
 *        7 = ghost-force exchange (pp_md_exchange_device) */
double pp_last_phase_ms(pp_context* ctx, int which);
void* pp_stream(pp_context* ctx);

#ifdef __cplusplus
}
#endif
#endif /* PROPHET_B200_H */
