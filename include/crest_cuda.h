/* crest_cuda.h — C ABI of the B200-native CREGEN duplicate / rotamer sorting path.
 *
 * Drop-in boundary: the replacement sits INSIDE the reference's
 *   cregen_CRE            /root/reference/src/cregen.f90:1125-1506
 * taking over K0-K4 (lines 1222-1376: axis, pair RMSD loops, duplicate sweep) and K5-K6
 * (lines 1421-1469: enuc, rotamer grouping).  Everything crossing the boundary is a plain
 * pointer or scalar so that Fortran can bind it with iso_c_binding exactly the way the reference
 * binds its own C code (src/symmetry2.f90:23-39 -> src/symmetry_i.c:1767).
 *
 * Conventions
 *   - arrays are caller-owned, column-major as Fortran passes them: xyz(3,nat,nall) in Angstrom
 *   - indices stored in dup[] / group[] are 1-based, 0 = "none" (as the reference's double()/group())
 *   - energies er[] in Hartree, ASCENDING (the post-condition of cregen_esort, cregen.f90:1038-1121)
 *   - every function returns 0 on success, non-zero on error; text via crest_cuda_last_error()
 *   - there is NO CPU fallback: without a CUDA device every compute entry point fails
 *   - not re-entrant: call from one host thread (the reference calls from its serial region)
 */
#ifndef CREST_CUDA_H
#define CREST_CUDA_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct crest_cuda_ctx crest_cuda_ctx;

/* ---- thresholds of type(systemdata) that cross the boundary (src/classes.f90:332-345) ---- */
typedef struct crest_cuda_thresholds {
  double ethr;      /* env%ethr   kcal/mol, energy gate (cregen.f90:1275,1356,1444) */
  double rthr;      /* env%rthr   Angstrom, RMSD threshold (cregen.f90:1364,1367)   */
  double bthr;      /* env%bthr2  rot.-const. threshold (cregen.f90:1368 uses 0.5*bthr, 1446 bthr) */
  double bthrmax;   /* env%bthrmax   (rotcompare.f90:127-142) */
  double bthrshift; /* env%bthrshift (rotcompare.f90:127-142) */
} crest_cuda_thresholds;

/* ------------------------------------------------------------------------------------------
 * Engine: the process-global set of GPUs the one-call entry points use.  CREST is one
 * process with OpenMP and no MPI (newcregen, cregen.f90:221-235, calls from the serial region), so ONE
 * host thread drives 1-8 GPUs: per-device contexts and streams, xyz split over the GPUs' PCIe links and
 * exchanged over NVLink peer copies, each GPU running its shard of the pair work, links combined by
 * system-scope atomicMax into GPU 0's dup[] over peer memory (no collective afterwards).
 *   ngpu_requested <= 0: every visible GPU; 1: the CURRENT device; n > 1: devices 0..n-1 (needs all-to-all
 *   peer access, else the engine stays on one GPU).  *ngpu_used receives the count in use.
 * Optional: without it the first one-call entry point initialises CREST_CUDA_NGPU GPUs (default 1).
 * crest_cuda_finalize releases every device buffer, stream and pinned staging buffer of the engine;
 * calling an entry point afterwards re-initialises it.
 * ------------------------------------------------------------------------------------------ */
int crest_cuda_init(int ngpu_requested, int* ngpu_used);
void crest_cuda_finalize(void);
int crest_cuda_ngpu(void);   /* GPUs of the engine, 0 before init / after finalize */

/* ------------------------------------------------------------------------------------------
 * One-call entry points (what the Fortran shim binds), on the engine's GPUs
 * ------------------------------------------------------------------------------------------ */

/* Replaces cregen.f90:1222-1376 (K0 axis + rot, K1-K3 pair RMSD incl. mirror, K4 sweep+backtrack).
 *   at(nat)            atomic numbers
 *   xyz(3,nat,nall)    input geometries (not modified)
 *   er(nall)           energies, ascending
 *   rmsd_mask(nat)     >0 = atom enters the RMSD (cregen.f90:1293-1299 heavy mask / includeRMSD);
 *                      NULL = all atoms
 *   rot_from_mask      != 0: rot. constants from the masked sub-structure (substruc,
 *                      cregen.f90:1225-1230); 0: from all atoms
 *   enantio            != 0: also compare the mirror image, keep the minimum (cregen.f90:1279-1284)
 * outputs
 *   rot(3,nall)        rot. constants in MHz (axis_module.f90:96-176)
 *   xyz_axis(3,nat,nall) geometries shifted to CMA and rotated into the principal-axis frame
 *                      (what axis() leaves in xyz, cregen.f90:1224); may be NULL
 *   dup(nall)          double(i) after backtrack: 0 = kept, else 1-based root it duplicates
 *   npairs             number of (i,j) pairs that passed the energy gate (= klong); may be NULL
 */
int crest_cuda_cregen_dup(int nat, int nall, const int* at, const double* xyz, const double* er,
                          const int* rmsd_mask, int rot_from_mask, int enantio,
                          const crest_cuda_thresholds* thr, double* rot, double* xyz_axis,
                          int* dup, long long* npairs);

/* Replaces cregen.f90:1421-1469 (K5 enuc, K6 greedy rotamer grouping, group ids) on the kept,
 * already partitioned ensemble.  group(0:nkept): group[0] = number of groups. enuc may be NULL. */
int crest_cuda_cregen_group(int nat, int nkept, const int* at, const double* xyz,
                            const double* er, const double* rot,
                            const crest_cuda_thresholds* thr, int* group, double* enuc);

/* K6 (group ids only) for callers that cannot call cudaMalloc: code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it.  NO counterpart in the reference.
 * Allocates nothing (no cudaMalloc / cudaFree / pool, no host allocation): inputs and the output
 * are DEVICE arrays of the caller on the current device - er(nkept) ascending, rot(3,nkept), enuc(nkept)
 * (K5's result, e.g. from crest_cuda_cregen_group), group(0:nkept) - and the launch is asynchronous on
 * `stream` (the caller synchronises).  The kernel's scratch (head flag + rotaniso per structure) is ONE
 * file-scope device array of 9 bytes x crest_cuda_noalloc_max_kept() = 32768 structures (288 KB, held by
 * every GPU that loads the library until exit); nkept beyond that is rejected with an error.  Every call
 * shares that array: calls that overlap on the GPU (two streams, two host threads) race - serialise them.
 * Same decisions as crest_cuda_cregen_group (the single-CTA greedy sweep of cregen.f90:1437-1469). */
int crest_cuda_cregen_group_noalloc(int nkept, const double* d_er, const double* d_rot, const double* d_enuc,
                                    const crest_cuda_thresholds* thr, int* d_group, void* stream);
int crest_cuda_noalloc_max_kept(void);

/* Whole cregen_CRE (cregen.f90:1195-1474) in one call, incl. the kept-first partition
 * (cregen.f90:1397-1419).  xyz is rotated AND permuted in place, er permuted in place.
 *   order(nall)  out: new 1-based position of input structure i (orderref); comments are permuted
 *                by the Fortran caller with it (character data never crosses the boundary)
 *   nallout      out: number of structures the caller keeps (cregen.f90:1393-1394,1470-1474)
 *   nosort       cregen.f90:1133 flag */
int crest_cuda_cregen_cre(int nat, int nall, const int* at, double* xyz, double* er,
                          const int* rmsd_mask, int rot_from_mask, int enantio, int nosort,
                          const crest_cuda_thresholds* thr, double* rot, int* order, int* group,
                          int* nallout, long long* npairs);

/* Diagnostics / parity: RMSD (ls_rmsd.f90:72-174, option 0) of explicit pairs of structures,
 * proper and mirror (x of the second structure negated, cregen.f90:1280).  packed(3,A,n).
 * i_idx/j_idx 0-based. */
int crest_cuda_rmsd_pairs(int A, int n, const double* packed, long long npairs, const int* i_idx,
                          const int* j_idx, double* rmsd, double* rmsd_mirror);

/* Rot. constants only: axis(nat,at,coord,rot,avmom) = axis_1 -> axis_0 (axis_module.f90:96-176). */
int crest_cuda_rotconst(int nat, int n, const int* at, const double* xyz, double* rot);

/* ---- rows either side of cregen_CRE ----
 * cregen_esort (cregen.f90:1038-1121) without the file/comment handling: sorts energies ascending
 * with the reference's own unstable quicksort (quicksort.f90:140-165; tie order is part of the
 * contract), permutes xyz(3,nat,nall) with it (xyz may be NULL), returns order(nall) = new 1-based
 * position of input structure i (what the caller permutes `comments` with) and nallout after the
 * energy-window cut `ewin` (kcal/mol).  Host code: O(N log N), serial in the reference too. */
int crest_cuda_cregen_esort(int nat, int nall, double* xyz, double* energies, double ewin,
                            int* order, int* nallout);

/* cregen_rmsdalign (cregen.f90:2229-2283): rotate structures 2..nall onto structure 1 by the
 * optimal heavy-atom (at > 2) rotation of rmsd() (ls_rmsd.f90:72-221, option 1), in place. */
int crest_cuda_rmsdalign(int nat, int nall, const int* at, double* xyz);

/* ---- ensemble files either side of CREGEN (host code, src/strucreader.f90) ----
 * rdensembleparam (268-309): number of readable structures and the largest atom count of an
 * xyz ensemble / trajectory file; a missing file gives nat = nall = 0 like the reference. */
int crest_cuda_rdensemble_param(const char* fname, int* nat, int* nall);

/* rdensemble_conf3 (433-481) + grepenergy (2225-2260): reads up to nall structures of nat atoms
 * (structures with another atom count are skipped, ensemble_strucskip) into at(nat),
 * xyz(3,nat,nall) and er(nall) = grepenergy(comment line); comments (nall records of comment_len
 * blank-padded characters = the reference's character(len=*) array) may be NULL, so may er and
 * nread (number of structures actually read).  Coordinate lines follow coordline (2005-2022):
 * "x y z sym" or "sym x y z"; labels map through e2i (2120-2147).  An unreadable coordinate line
 * or a truncated last structure is an error (the reference: `error stop` on the latter). */
int crest_cuda_rdensemble(const char* fname, int nat, int nall, int* at, double* xyz, double* er,
                          char* comments, int comment_len, int* nread);

/* wrxyz_channel (1679-1698) for every structure, as the ensemble writers of cregen.f90:2143-2151
 * do: '(2x,i0)' nat, the trimmed comment record, '(1x,a2,1x,3f20.10)' per atom with
 * i2e(at,'nc').  comments == NULL writes the energy er(i) alone (blank line if er is NULL too). */
int crest_cuda_wrensemble(const char* fname, int nat, int nall, const int* at, const double* xyz,
                          const double* er, const char* comments, int comment_len);

/* grepenergy (2225-2260) and e2i (2120-2147) on their own. */
double crest_cuda_grepenergy(const char* line);
int crest_cuda_e2i(const char* label);

/* ---- callers of rmsd() beside cregen_CRE ----
 * compare_ensembles main loop (src/ensemblecomp.f90:213-248): rmat(i,j) = minimum all-atom RMSD
 * (ls_rmsd.f90:72-174, no mirror) between any rotamer of conformer i of ensemble 1 and any rotamer
 * of conformer j of ensemble 2 (the reference's seed `rmat2 = 100` never survives: every rotamer
 * pair overwrites its slot before minval()).
 *   xyz1(3,nat,nall1), xyz2(3,nat,nall2)   the two ensembles as rdensemble delivers them
 *   b1/e1(nconf1), b2/e2(nconf2)           1-based inclusive rotamer ranges from the .cretrack
 *                                          files (ensemblecomp.f90:123-126,161-164)
 *   rmat(nconf1,nconf2)                    out, column-major double */
int crest_cuda_compare_blocks(int nat, int nall1, const double* xyz1, int nconf1, const int* b1,
                              const int* e1, int nall2, const double* xyz2, int nconf2,
                              const int* b2, const int* e2, double* rmat);

/* "Correlation between Conformers" walk (src/ensemblecomp.f90:258-300) as data instead of print
 * lines: line n pairs left[n] with right[n] (1-based conformer numbers, 0 = blank column).
 * en1(nconf1), en2(nconf2) energies of the lowest rotamers; left/right hold nconf1+nconf2 entries.
 * Host code.  Returns the number of lines, -1 on a bad argument. */
int crest_cuda_compare_correlate(int nconf1, int nconf2, const double* en1, const double* en2,
                                 const double* rmat, double rthr, int* left, int* right);

/* intraconfRMSD (src/entropy/entropic.f90:938-1097): for every conformer with nrot > 1, RMSDs
 * between its rotamers low(c)..low(c)+nrot(c)-1 (1-based) on the atoms with inc(k) == 1;
 * fac(c) = uniqueCore(rmat, nrot, 0.125) (1074-1095), enantiofac(c) = 2 when some pair has
 * rmsd > 0.125 and mirror rmsd < 0.09375 (1024-1031), then spread over equal symmetry symbols
 * symsym (3 characters per conformer, blank padded; symmap 1096-1116).  enantiofac is in/out
 * (the caller seeds it with 1, entropic.f90:137).  sum(inc) < 2 sets fac = 1 and returns. */
int crest_cuda_intraconf(int nat, int nall, const double* xyz, const int* inc, int nconf,
                         const int* low, const int* nrot, const char* symsym, double* fac,
                         double* enantiofac);

/* De-duplication inside the genetic structure crossing (src/algos/crossing.f90:227-338,
 * the part between GMETRY / the CN clash check and the output): cand(3,nat,ncand) are the generated
 * structures in the reference's single-thread loop order (i, then j < i), xyzref(3,nat) the lowest
 * conformer, rthr in the unit of the coordinates (the reference works in Bohr).  Per structure:
 * rval = rmsd(cand, xyzref) (263); rval <= rthr discards it; otherwise it is compared with the accepted
 * structures whose own rval lies within rthr/100 (284) and rejected if any RMSD is < rthr (287-291), else
 * appended, or - with the maxgen2 slots full - ends the search (huge_number != 0, 303-304) or replaces the
 * slot sdselect picks (306-310, 401-432).  The GPU evaluates rval and every gated pair RMSD; the ordered
 * replay of the critical section is host code.
 *   taken(maxgen2)  out: 1-based index into cand of the structure in accepted slot k, 0 = empty
 *   rms(maxgen2)    out: its rval;   ntaken out: slots in use
 *   counters(4)     out: [0] discarded by rval <= rthr, [1] ident (duplicates of accepted structures),
 *                   [2] rcount, [3] structures looked at before the search stopped
 *   rmsdsum         out: sum of rval over the rcount structures (rmsdavg before the division, 340) */
int crest_cuda_crossing_dedup(int nat, int ncand, const double* cand, const double* xyzref, double rthr,
                              int maxgen2, int huge_number, int* taken, double* rms, int* ntaken,
                              long long* counters, double* rmsdsum);

/* All-pairs float32 RMSD band exactly as the reference's rmat (cregen.f90:1260-1349): row i holds
 * j = 0..i-1 at offset i*(i-1)/2 (ethr = infinity layout).  For tests on small ensembles. */
int crest_cuda_rmat_allpairs(int nat, int nall, const int* at, const double* xyz,
                             const int* rmsd_mask, int enantio, float* rmat);

/* ------------------------------------------------------------------------------------------
 * Context API: device-resident ensemble, explicit stream, tile-space sharding for multi-GPU.
 * One process per GPU; the caller owns the collective (NCCL all-reduce MAX on the dup buffer).
 * ------------------------------------------------------------------------------------------ */
int crest_cuda_ctx_create(crest_cuda_ctx** ctx, int device /* -1 = current */);
void crest_cuda_ctx_destroy(crest_cuda_ctx* ctx);

/* Copy the ensemble host->device and keep it resident. Does not launch math.  xyz may be NULL
 * (reserve only, see crest_cuda_ctx_upload_xyz_rows). */
int crest_cuda_ctx_upload(crest_cuda_ctx* ctx, int nat, int nall, const int* at,
                          const double* xyz, const double* er, const int* rmsd_mask,
                          int rot_from_mask);

/* Multi-GPU upload: after crest_cuda_ctx_upload(..., xyz = NULL, ...) each rank copies only ITS rows
 * [first, first+count) of xyz(3,nat,nall) host->device; the caller then all-gathers the slices
 * over NVLink into the buffer crest_cuda_ctx_xyz_devptr returns (capacity 3*nat*(nall+64) doubles,
 * so equal slices of ceil(nall/ranks) rows fit for up to 64 ranks). */
int crest_cuda_ctx_upload_xyz_rows(crest_cuda_ctx* ctx, int first, int count, const double* xyz_rows);
void* crest_cuda_ctx_xyz_devptr(crest_cuda_ctx* ctx);

/* K0 on the device-resident ensemble: CMA shift, principal axes, rot, aniso, pack (kernel 1). */
int crest_cuda_ctx_pack(crest_cuda_ctx* ctx, void* stream);

/* Kernels (2)+(3): pair tiles of shard `rank` of `nranks` (contiguous slice of the in-band tile
 * list).  Accumulates into the context's device dup buffer with max (zeroed by ctx_pack).
 * want_mask != 0 additionally writes the packed 8x8-tile duplicate bitmask. */
int crest_cuda_ctx_pairs(crest_cuda_ctx* ctx, int enantio, const crest_cuda_thresholds* thr,
                         int rank, int nranks, int want_mask, void* stream);

/* Device pointer (int32[nall], 1-based parents, 0 = none) for the caller's all-reduce(MAX). */
void* crest_cuda_ctx_dup_devptr(crest_cuda_ctx* ctx);
/* Device pointer of the pair counter (int64[1]) for the caller's all-reduce(SUM). */
void* crest_cuda_ctx_npairs_devptr(crest_cuda_ctx* ctx);

/* Kernel (4): backtrack (root chase) on the device dup buffer. */
int crest_cuda_ctx_sweep(crest_cuda_ctx* ctx, void* stream);

/* Results device->host. Any pointer may be NULL. Synchronises the stream. */
int crest_cuda_ctx_download(crest_cuda_ctx* ctx, double* rot, double* xyz_axis, int* dup,
                            long long* npairs, void* stream);

/* Geometry of the launch that ctx_pairs would do (for bench/roofline bookkeeping):
 * info[0]=work items of this rank, [1]=work items total (FP64-only kernels; = [0] on the screen path), [2]=atoms in RMSD (A), [3]=padded atoms used by MMA,
 * [4]=grid, [5]=dynamic smem bytes, [6]=kernels launched by the last ctx_* call sequence,
 * [7]=pair-kernel variant of the last ctx_pairs (0 row-tile DMMA, 1 column DMMA, 2 tensor-core
 * TF32 screen + exact FP64 candidate scan), [8]=candidate pairs the screen could not prove far,
 * [9]=work items the classifier found candidate-dense, [10]=screen violations counted in the
 * CREST_CUDA_DEBUG=3 self-check (must be 0), [11]=exact FP64 evaluations the top-down scan spent
 * (<= [8]: a row ends at its first match), [12]=flags: 1 unsafe magnitudes (every item re-done in FP64),
 * 2 dense work list overflowed (items left to the scan), [13]=rows that met a candidate-dense item,
 * [14]=those that ended inside their evaluation budget.  Counters accumulate since the last ctx_pack.
 * info must hold 16 entries; reading [8..12] synchronises the device. */
int crest_cuda_ctx_info(crest_cuda_ctx* ctx, const crest_cuda_thresholds* thr, int rank,
                        int nranks, long long* info);

/* Evidence for the screen's error budget: after ctx_upload + ctx_pack, runs the tensor-core
 * screen over ALL pairs with the raw TMEM covariances dumped (9*nall^2 floats of device memory) and
 * recomputes every dot product of the same TF32-rounded operands exactly: max_rel_err = the largest
 * |R~_ab - R_ab| / (|x_a||y_b|) the tcgen05 FP32 accumulation produced, ndots = dot products checked,
 * gamma_budget = what the bound allows for this K extent, kappa = the resulting test-point lowering. */
int crest_cuda_ctx_tf32_accum(crest_cuda_ctx* ctx, double* max_rel_err, long long* ndots,
                              double* gamma_budget, double* kappa);

/* Copy of the device duplicate bitmask written by ctx_pairs(want_mask=1): n8*n8 64-bit words with
 * n8 = ceil(nall/8); bit r*8+c of word (ti,tj) = pair (i=8*ti+r, j=8*tj+c) is a duplicate link. */
int crest_cuda_ctx_download_mask(crest_cuda_ctx* ctx, unsigned long long* mask, void* stream);

/* ---- host-only bookkeeping: no device is touched (used by the multi-rank driver and tests) ----
 * e_threshold: the double E with  e < E  <=>  real(sqrt(max(0,e)/n),sp) < thr, i.e. the
 *   reference's test of the float32-stored RMSD against rthr / 2*rthr (cregen.f90:1176,1286,
 *   1359-1367) moved onto e = |x|^2+|y|^2-2*lambda so the kernel needs no sqrt/div per pair.
 * band_tiles: in-band tile list of the pair triangle for ascending energies (the reference's
 *   rmap1/rmap2 band, cregen.f90:1240-1256, at tile granularity): i-panel I owns the j-tiles
 *   jfirst[I] .. jfirst[I]+(rowstart[I+1]-rowstart[I])-1.  rowstart has ceil(nall/tile_i)+1
 *   entries, jfirst ceil(nall/tile_i).
 * shard_range: the contiguous slice [begin,end) of the tile list that `rank` of `nranks` owns. */
double crest_cuda_e_threshold(double thr, int n);
void crest_cuda_tile_shape(int* tile_i, int* tile_j);
int crest_cuda_band_tiles(int nall, const double* er, double ethr, long long* rowstart,
                          int* jfirst, long long* ntiles);
int crest_cuda_shard_range(long long ntiles, int rank, int nranks, long long* begin,
                           long long* end);
/* Work list of the tensor-core screen kernel (host-only): quadruples {I, g_begin, g_end,
 * g_gate_end}: i-tile I = conformers 128*I .. 128*I+127 against the j-groups g (16 conformers
 * each) in [g_begin, g_end); groups below g_gate_end hold pairs that can fail the energy gate
 * (cregen.f90:1275) and are gated pair by pair.  items may be NULL to query the count; cap is the
 * capacity of items in quadruples. */
int crest_cuda_tf32_items(int nall, const double* er, double ethr, int* items, long long cap,
                          long long* nitems);
/* Multi-rank: the pair space is shared by ROWS - rank r owns the i-tiles I with I % nranks == r and all
 * their columns (the per-row "largest j" then needs no exchange beyond combining disjoint rows, and links
 * found in a row's high columns prune its low ones on the same rank).  tf32_items_rank = the list of one
 * rank (items cut so that it still has >= 16 per CTA); tf32_items_n = all rank lists concatenated. */
int crest_cuda_tf32_items_rank(int nall, const double* er, double ethr, int nranks, int rank, int* items,
                               long long cap, long long* nitems);
int crest_cuda_tf32_items_n(int nall, const double* er, double ethr, int nranks, int* items,
                            long long cap, long long* nitems);

/* Developer knobs (same syntax as the CREST_CUDA_TUNE environment variable; NULL = defaults):
 * "notf32" forces the FP64 (DMMA) pair kernels, "tile" the row-tile kernel, "chunk=<n>" j-groups per
 * work item, "maskmb=<MB>" the candidate-bitmask budget, "forcedense"/"nodense"/"densebudget=<n>" the
 * dense-item policy, "scanocc=<2|3|4>" resident blocks of the exact scan, "packold" / "packg=-<G>" the
 * multi-sweep pack kernel / the lanes per conformer of the one-pass one, "blocks=<n>" equal row blocks of the
 * upload pipeline, "evenblocks", "blockf1=<%>", "blockgain=<%>" its schedule, "nostage" no pinned staging of
 * the whole-CREGEN geometry, "copythr=<n>" staging threads, "seqgroup" / "groupcap=<n>" the K6 path.
 * Takes effect at the next ctx_upload. */
void crest_cuda_set_tune(const char* spec);

const char* crest_cuda_last_error(void);
int crest_cuda_device_count(void);

#ifdef __cplusplus
}
#endif
#endif /* CREST_CUDA_H */
