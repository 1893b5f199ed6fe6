/* libteqpb200 -- C ABI of the B200-native batched equation-of-state evaluator.
 *
 * Part 1 is the reference's extern "C" interface, exported with identical names, argument order, ownership and error
 * conventions (/root/reference/interface/C/teqpc.h:2-20, implementation /root/reference/interface/C/teqpc.cpp:75-225), so that
 * existing ctypes / FFI callers (e.g. /root/reference/interface/C/test_teqpc.py) bind to this library unchanged.  Scalar
 * calls are evaluated ON THE GPU as a batch of one state point: there is no CPU evaluation path in this library.
 *
 * Part 2 adds the batched entry points (proposal in SURVEY.md section 8b): structure-of-arrays buffers owned by the caller,
 * int return code (0 = OK), message copied into errmsg with the reference's truncation rules.
 *
 * Conventions for part 2
 *   - T[npts], rho[npts]; molefrac[Ncomp][npts] (row i = component i, row stride npts), or molefrac[Ncomp] for every point
 *     when TEQPB200_BROADCAST_X is set; rhovec[Ncomp][npts] for the isochoric family.
 *   - outputs are [rows][npts] with row stride npts.
 *   - flags & TEQPB200_DEVICE_PTRS: all data pointers are device pointers on the CURRENT CUDA device; the call enqueues on
 *     the stream set with teqpb200_set_stream() (default: the legacy default stream) and returns without synchronising.
 *     Otherwise the pointers are host pointers (pinned memory from teqpb200_host_alloc() makes the copies asynchronous);
 *     the batch is split evenly over the devices chosen with teqpb200_set_devices(), streamed through each device in
 *     chunks with copies overlapped with compute, and the call returns when all results are in the output buffers.
 */
#ifndef TEQPB200_H
#define TEQPB200_H

#ifdef __cplusplus
extern "C" {
#endif

/* ---------------- Part 1: the reference C ABI (teqpc.h:2-20) ---------------- */
int build_model(const char* j, long long int* uuid, char* errmsg, int errmsg_length);                       /* teqpc.h:2  */
int free_model(const long long int uuid, char* errmsg, int errmsg_length);                                  /* teqpc.h:4  */
int get_Arxy(const long long int uuid, const int NT, const int ND, const double T, const double rho,
             const double* molefrac, const int Ncomp, double* val, char* errmsg, int errmsg_length);         /* teqpc.h:6  */
int get_ATrhoXi(const long long int uuid, const double T, const int NT, const double rhomolar, const int ND,
                const double* molefrac, const int Ncomp, const int i, const int NXi, double* val,
                char* errmsg, int errmsg_length);                                                            /* teqpc.h:8  */
int get_ATrhoXiXj(const long long int uuid, const double T, const int NT, const double rhomolar, const int ND,
                  const double* molefrac, const int Ncomp, const int i, const int NXi, const int j, const int NXj,
                  double* val, char* errmsg, int errmsg_length);                                             /* teqpc.h:10 */
int get_ATrhoXiXjXk(const long long int uuid, const double T, const int NT, const double rhomolar, const int ND,
                    const double* molefrac, const int Ncomp, const int i, const int NXi, const int j, const int NXj,
                    const int k, const int NXk, double* val, char* errmsg, int errmsg_length);              /* teqpc.h:12 */
int get_AtaudeltaXi(const long long int uuid, const double tau, const int Ntau, const double delta, const int Ndelta,
                    const double* molefrac, const int Ncomp, const int i, const int NXi, double* val,
                    char* errmsg, int errmsg_length);                                                        /* teqpc.h:14 */
int get_AtaudeltaXiXj(const long long int uuid, const double tau, const int Ntau, const double delta, const int Ndelta,
                      const double* molefrac, const int Ncomp, const int i, const int NXi, const int j, const int NXj,
                      double* val, char* errmsg, int errmsg_length);                                         /* teqpc.h:16 */
int get_AtaudeltaXiXjXk(const long long int uuid, const double tau, const int Ntau, const double delta, const int Ndelta,
                        const double* molefrac, const int Ncomp, const int i, const int NXi, const int j, const int NXj,
                        const int k, const int NXk, double* val, char* errmsg, int errmsg_length);          /* teqpc.h:18 */
int get_dmBnvirdTm(const long long int uuid, const int Nvir, const int NT, const double T, const double* molefrac,
                   const int Ncomp, double* val, char* errmsg, int errmsg_length);                           /* teqpc.h:20 */

/* ---------------- Part 2: batched entry points ---------------- */
#define TEQPB200_DEVICE_PTRS 1
#define TEQPB200_BROADCAST_X 2

/* AbstractModel::get_Ar0{nmax}n (teqpcpp.hpp:111-113; derivs.hpp:543-569,621-624): out[n][k] = rho^n d^n alphar/drho^n, n = 0..nmax (nmax 0..6) */
int get_Ar0n_batch(const long long int uuid, const int nmax, const long long int npts, const double* T, const double* rho,
                   const double* molefrac, const int Ncomp, double* out, const int flags, char* errmsg, int errmsg_length);
/* AbstractModel::get_Ar{nmax}0n (teqpcpp.hpp:115-117; derivs.hpp:571-611): out[n][k] = (1/T)^n d^n alphar/d(1/T)^n, n = 0..nmax (nmax 0..4) */
int get_Arn0_batch(const long long int uuid, const int nmax, const long long int npts, const double* T, const double* rho,
                   const double* molefrac, const int Ncomp, double* out, const int flags, char* errmsg, int errmsg_length);
/* AbstractModel::get_Arxy for several (NT[p], ND[p]) at once (teqpcpp.hpp:104-109; derivs.hpp:150-238): out[p][k].  Reachable pairs, the same
   for every model family: NT = 0 with ND <= 6; ND = 0 with NT <= 4; NT <= 2 with ND <= 4 (a superset of the reference's runtime dispatch,
   NT + ND <= 3).  The pairs are partitioned over at most three sweeps (density row, temperature column, mixed block), each run once. */
int get_Arxy_batch(const long long int uuid, const int npairs, const int* NT, const int* ND, const long long int npts,
                   const double* T, const double* rho, const double* molefrac, const int Ncomp, double* out,
                   const int flags, char* errmsg, int errmsg_length);
/* every get_ATrhoXi combination in one call (derivs.hpp:247-292): out[((i*3+NT)*3+ND)*4+NXi][k], NT,ND in 0..2, NXi in 0..3 */
int get_ATrhoXi_table_batch(const long long int uuid, const long long int npts, const double* T, const double* rho,
                            const double* molefrac, const int Ncomp, double* out, const int flags,
                            char* errmsg, int errmsg_length);
/* cross-composition derivatives of one component pair (i, j) in one sweep (derivs.hpp:294-314, 423-442; replaces the per-combination
   calls of teqpc.h:8): out[((NT*2+ND)*3+NXi)*3+NXj][k], NT, ND in 0..1, NXi, NXj in 0..2 (x_i and x_j independent) */
int get_ATrhoXiXj_table_batch(const long long int uuid, const int i, const int j, const long long int npts, const double* T,
                              const double* rho, const double* molefrac, const int Ncomp, double* out, const int flags,
                              char* errmsg, int errmsg_length);
/* ... and of one component triple (derivs.hpp:316-339, 444-465; teqpc.h:9): out[(((NT*2+ND)*2+NXi)*2+NXj)*2+NXk][k], all orders 0..1 */
int get_ATrhoXiXjXk_table_batch(const long long int uuid, const int i, const int j, const int k, const long long int npts,
                                const double* T, const double* rho, const double* molefrac, const int Ncomp, double* out,
                                const int flags, char* errmsg, int errmsg_length);
/* IsochoricDerivatives::build_Psir_fgradHessian_autodiff (derivs.hpp:1051-1069): f[k], g[i][k], H[i*Ncomp+j][k]; any of f,g,H may be NULL */
int build_Psir_fgradHessian_batch(const long long int uuid, const long long int npts, const double* T, const double* rhovec,
                                  const int Ncomp, double* f, double* g, double* H, const int flags,
                                  char* errmsg, int errmsg_length);
/* The derived isochoric properties of AbstractModel (teqpcpp.hpp:52-75, 128-160; derivs.hpp:962-978, 1196-1201, 1264-1437, 1482-1494) in one
   launch; every output may be NULL.  Scalars [k]: get_pr, get_splus, get_dpdT_constrhovec.  Vectors [i][k]: build_d2PsirdTdrhoi_autodiff,
   get_dpdrhovec_constT, get_partial_molar_volumes, get_chempotVLE_autodiff, get_dchempotdT_autodiff,
   get_d_ln_fugacity_coefficients_{dT_constrhovec, drho_constTmolefracs, dv_constTmolefracs}.  Matrix [i*Ncomp+j][k]:
   get_d_ln_fugacity_coefficients_dmolefracs_constTrho (d ln phi_i / d x_j; not symmetric). */
int isochoric_properties_batch(const long long int uuid, const long long int npts, const double* T, const double* rhovec, const int Ncomp,
                               double* pr, double* splus, double* dpdT, double* d2PsirdTdrhoi, double* dpdrhovec, double* pmv,
                               double* chempotVLE, double* dchempotdT, double* dlnphi_dT, double* dlnphi_drho, double* dlnphi_dv,
                               double* dlnphi_dx, const int flags, char* errmsg, int errmsg_length);
/* build_Psir_fgradHessian_batch AND isochoric_properties_batch in one call (BASELINE config C5: "isochoric Hessian / fugacity composition derivatives"):
   the same outputs, bit for bit; for the GERG / multifluid term-program models ONE launch of the lane-cooperative kernel produces both (the passes over
   the reducing pairs and the term program are shared), other families run their two launches inside the call.  Every output may be NULL. */
int build_Psir_fgradHessian_and_properties_batch(const long long int uuid, const long long int npts, const double* T, const double* rhovec, const int Ncomp,
                                                 double* f, double* g, double* H, double* pr, double* splus, double* dpdT, double* d2PsirdTdrhoi,
                                                 double* dpdrhovec, double* pmv, double* chempotVLE, double* dchempotdT, double* dlnphi_dT, double* dlnphi_drho,
                                                 double* dlnphi_dv, double* dlnphi_dx, const int flags, char* errmsg, int errmsg_length);
/* IsochoricDerivatives::get_ln_fugacity_coefficients / get_fugacity_coefficients (derivs.hpp:1208-1230): out[i][k] */
int get_ln_fugacity_coefficients_batch(const long long int uuid, const long long int npts, const double* T, const double* rhovec,
                                       const int Ncomp, double* out, const int flags, char* errmsg, int errmsg_length);
int get_fugacity_coefficients_batch(const long long int uuid, const long long int npts, const double* T, const double* rhovec,
                                    const int Ncomp, double* out, const int flags, char* errmsg, int errmsg_length);

/* every get_AtaudeltaXi combination in one call (derivs.hpp:341-351; models with alphar_taudelta only, i.e. kind "multifluid"): layout of
   get_ATrhoXi_table_batch with tau^NT delta^ND d^(NT+ND+NXi) alphar / dtau^NT ddelta^ND dx_i^NXi; host pointers */
int get_AtaudeltaXi_table_batch(const long long int uuid, const long long int npts, const double* tau, const double* delta, const double* molefrac,
                                const int Ncomp, double* out, const int flags, char* errmsg, int errmsg_length);
/* VirialDerivatives::get_dmBnvirdTm (teqpcpp.hpp:131; derivs.hpp:780-841) for many temperatures: out[k] = d^NT B_Nvir / dT^NT.  Nvir 2..7 for
   NT = 0, 2..5 for NT = 1, 2 (the scalar reference entry above serves Nvir 2..4); composition rows as in get_Ar0n_batch */
int get_dmBnvirdTm_batch(const long long int uuid, const int Nvir, const int NT, const long long int npts, const double* T, const double* molefrac,
                         const int Ncomp, double* out, const int flags, char* errmsg, int errmsg_length);
/* IsochoricDerivatives::get_Psir_sigma_derivs (teqpcpp.hpp:160; derivs.hpp:1496-1509): out[k][p] = d^k Psir(T, rhovec + sigma v) / dsigma^k at
   sigma = 0, k = 0..4; rhovec and v are [Ncomp][npts] */
int get_Psir_sigma_derivs_batch(const long long int uuid, const long long int npts, const double* T, const double* rhovec, const double* v,
                                const int Ncomp, double* out, const int flags, char* errmsg, int errmsg_length);

/* Batched caller of the hot path: teqp::pure_VLE_T / AbstractModel::pure_VLE_T
   (/root/reference/include/teqp/algorithms/VLE_pure.hpp:18-143; teqpcpp.hpp) for npts isotherms at once -- Newton's method on
   p(T, rhoL) = p(T, rhoV), mu(T, rhoL) = mu(T, rhoV) from the guesses rhoL0[k], rhoV0[k], at most maxiter steps with the reference's stopping
   rules per point, every iteration ONE launch of the Lambda table kernel over the 2 npts liquid + vapour densities and one update kernel, no
   host round trip.  molefrac[Ncomp] is ONE composition for all points ([1.0] for a one-component model; a pure fluid inside a mixture model is
   its unit vector), whatever the TEQPB200_BROADCAST_X flag says.  Optional outputs (NULL to skip): psat[k] = rhoV R T (1 + Ar01(rhoV)) and
   dpsatdT[k] = teqp::dpsatdT_pure (VLE_pure.hpp:158-171, Clapeyron).  A point that does not converge returns whatever the iteration produced
   (possibly the trivial solution rhoL = rhoV or non-finite values), as the reference does. */
int pure_VLE_T_batch(const long long int uuid, const long long int npts, const double* T, const double* rhoL0, const double* rhoV0, const int maxiter,
                     const double* molefrac, const int Ncomp, double* rhoL, double* rhoV, double* psat, double* dpsatdT, const int flags,
                     char* errmsg, int errmsg_length);

/* teqp::mix_VLE_Tx / AbstractModel::mix_VLE_Tx (/root/reference/include/teqp/algorithms/VLE.hpp:76-192) for npts binary-mixture problems at once: T[k]
   and the liquid mole fraction xspec[0][k] given, unknowns the molar concentrations of both phases; Newton's method on equal chemical potentials,
   equal pressure and the composition constraint, Jacobian from the Hessians of Psi^r, with the reference's step limiter, frozen zero-mole-fraction
   components and stopping rules per point.  Every iteration is ONE launch of the isochoric kernel (Psi^r, gradient, Hessian) over the 2 npts liquid +
   vapour states and one update kernel.  rhovecL0, rhovecV0, xspec, rhovecL, rhovecV: [2][npts].  return_code[k] (optional, doubles) holds
   teqp::VLE_return_code (VLE_types.hpp:44): 1 xtol_satisfied, 2 functol_satisfied, 4 maxiter_met, 5 notfinite_step, 0 when maxiter = 0.  Two-component
   models with a composition-independent gas constant only. */
int mix_VLE_Tx_batch(const long long int uuid, const long long int npts, const double* T, const double* rhovecL0, const double* rhovecV0,
                     const double* xspec, const double atol, const double reltol, const double axtol, const double relxtol, const int maxiter,
                     double* rhovecL, double* rhovecV, double* return_code, const int flags, char* errmsg, int errmsg_length);

/* teqp::mix_VLE_Tp (VLE.hpp:196-380, the Newton branch :351-367) for npts binary problems at once: T[k] and p[k] given, unknowns the molar concentrations of
   both phases; residuals equal chemical potentials, (p_V - p)/p and (p_L - p)/p; exactly maxiter Newton steps with the reference's step limiter and no stopping
   rule (the reference has none).  Launch structure and array layouts of mix_VLE_Tx_batch. */
int mix_VLE_Tp_batch(const long long int uuid, const long long int npts, const double* T, const double* p, const double* rhovecL0, const double* rhovecV0,
                     const int maxiter, double* rhovecL, double* rhovecV, const int flags, char* errmsg, int errmsg_length);

/* Per-point status of a finished batched call (the "per-point int* status" of SURVEY.md section 8b; the reference throws per call --
   teqp::InvalidArgument / IterationFailure, /root/reference/include/teqp/exceptions.hpp:17-64 -- where a batch can only flag the point):
   status[k] is a bit set computed on the GPU from the inputs of point k and the nrows output rows out[r][k] a previous entry point wrote for it
   (nrows = 0: inputs only).  Failures inside the models (e.g. the Barker-Henderson diameter iteration of SAFT-VR-Mie, saftvrmie.hpp d > sigma / j < 1)
   surface as non-finite outputs.  Same pointer / flag conventions as get_Ar0n_batch; status is int[npts] (a device pointer with TEQPB200_DEVICE_PTRS). */
#define TEQPB200_STATUS_BAD_INPUT 1          /* T <= 0, rho < 0, a mole fraction < 0, or any of them not finite */
#define TEQPB200_STATUS_NONFINITE_OUTPUT 2   /* an output row holds NaN / Inf at this point */
#define TEQPB200_STATUS_COMPOSITION_SUM 4    /* |sum(x) - 1| > 1e-6 (informational: the reference does not normalise or check) */
int teqpb200_point_status_batch(const long long int uuid, const long long int npts, const double* T, const double* rho, const double* molefrac,
                                const int Ncomp, const int nrows, const double* out, int* status, const int flags, char* errmsg, int errmsg_length);

/* ---------------- utilities ---------------- */
/* AbstractModel::get_reducing_temperature (which = 0) / get_reducing_density (which = 1), teqpcpp.hpp:124-125: multi-fluid families only */
int teqpb200_get_reducing(const long long int uuid, const int which, const double* molefrac, const int Ncomp, double* val, char* errmsg, int errmsg_length);
int teqpb200_get_R(const long long int uuid, const double* molefrac, const int Ncomp, double* R, char* errmsg, int errmsg_length); /* AbstractModel::get_R, teqpcpp.hpp:102 */
int teqpb200_get_ncomp(const long long int uuid, int* Ncomp, char* errmsg, int errmsg_length);
int teqpb200_device_count(void);
int teqpb200_set_devices(const int* device_ids, const int n, char* errmsg, int errmsg_length); /* host-pointer calls shard over these devices */
int teqpb200_set_stream(void* cuda_stream);                 /* stream for TEQPB200_DEVICE_PTRS calls made by this thread */
void* teqpb200_host_alloc(unsigned long long nbytes);       /* pinned host memory */
void teqpb200_host_free(void* p);
long long int teqpb200_kernel_launches(void);               /* kernels launched by this library so far (all threads) */
int teqpb200_release_scratch(void);                         /* frees the device staging arenas of the host-pointer path */
/* ---- no-allocation variant of get_Arxy_batch.  NO COUNTERPART IN THE REFERENCE: added for callers
   that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arguments, results and pair rules as
   get_Arxy_batch, TEQPB200_DEVICE_PTRS only (stream of teqpb200_set_stream), npairs <= 32.  The covering derivative table is kept in a static
   device array of teqpb200_noalloc_max_doubles() = 4 Mi doubles (32 MiB per GPU that loads the library, held until exit) instead of a
   cudaMallocAsync block: a call needing more than that many (scratch rows x npts) doubles is rejected with an error (code 1) -- split the batch.
   The call makes no cudaMalloc / cudaFree / cudaMallocAsync and calls no allocating library routine; for families whose kernels are generated
   per model (GERG, multifluid) the shape must have been used once through get_Arxy_batch / get_Ar0n_batch before (otherwise: error, not a
   compile).  All calls share the one array: calls that overlap on the GPU (other streams, other host threads) race on it.
   teqpb200_noalloc_setup() resolves the array's device address for the CURRENT device (once per device; the entry point does it at first use
   otherwise). */
int get_Arxy_batch_noalloc(const long long int uuid, const int npairs, const int* NT, const int* ND, const long long int npts,
                           const double* T, const double* rho, const double* molefrac, const int Ncomp, double* out, const int flags,
                           char* errmsg, int errmsg_length);
long long int teqpb200_noalloc_max_doubles(void);
int teqpb200_noalloc_setup(char* errmsg, int errmsg_length);
void teqpb200_inject_fault(const long long int chunk_index); /* TEST HOOK: the chunk_index-th staged chunk from now on fails (error-path tests) */
double teqpb200_last_call_device_ms(const int device);      /* device-side duration (first H2D .. last D2H, CUDA events) of the last
                                                               host-pointer call on that device; -1 if none */
int teqpb200_fp64_peak(double* tflops, char* errmsg, int errmsg_length); /* measured DFMA peak of the current device, TFLOP/s */
/* deterministic synthetic state points on the device: counter-hash uniforms u_j(k) of the global point
   index k0+k; T = Tlo + (Thi-Tlo) u0, rho log-uniform in [rholo, rhohi] from u1, x_i ~ z0_i exp(0.2 (2 u_{2+i} - 1)) renormalised.
   Device pointers on the current device; rows of x have stride npts. */
int teqpb200_synth_points(const long long int npts, const long long int k0, const unsigned long long seed, const double Tlo,
                          const double Thi, const double rholo, const double rhohi, const double* z0_host, const int Ncomp,
                          double* T, double* rho, double* x, char* errmsg, int errmsg_length);

/* Per-model generated kernels (GERG-2004/2008: csrc/tq_gerg_jit.hpp).  which = 0: the generated CUDA C translation unit for the
   Lambda^r table of shape (NT, ND), NUL-terminated; which = 1: its constant pool (doubles); which = 2: the sm_100a cubin NVRTC makes of
   it (no device needed).  Writes min(*nbytes, needed) bytes into buf (buf may be NULL) and returns the needed size in *nbytes. */
int teqpb200_generated_kernel(const long long int uuid, const int NT, const int ND, const int which, void* buf,
                              unsigned long long* nbytes, char* errmsg, int errmsg_length);

#ifdef __cplusplus
}
#endif
#endif
