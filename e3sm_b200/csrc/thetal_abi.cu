// C ABI of libthetal_b200.so: host-side session, device-resident state, step sequencing.
// Mirrors the HOMMEXX C interface the Fortran theta-l_kokkos driver binds
// (H/src/theta-l_kokkos/theta_f2c_mod.F90, impl cxx/cxx_f90_interface_theta.cpp,
//  H/src/share/cxx/prim_driver.cpp, mpi/mpi_cxx_f90_interface.cpp); see include/thetal_b200.h.
#include "../../include/thetal_b200.h"
#include "thetal_kernels.h"
#include "thetal_reprosum.h"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <array>
#include <cmath>
#include <map>
#include <set>
#include <algorithm>
#include <nccl.h>
#include <nvtx3/nvToolsExt.h>   // header-only; ranges carry the reference's GPTL timer names

using namespace thetal;

namespace {

struct Connection { int e1, e1_gid, dir1, e2_lid, e2_gid, e2_pid, dir2; };
struct Peer { int rank, send_off, send_cnt, recv_off, recv_cnt; };

struct Session {
  bool active = false;
  int rank = 0, nranks = 1, device = 0;
  int abort_on_error = 1;
  int last_error = 0;
  std::string last_msg;
  Params prm{};
  bool params_set = false;
  int nlev = 0;
  int nm1 = 0, n0 = 1, np1 = 2, nstep = 0, nstep0 = 0;   // 0-based
  int hypervis_order = 2;
  double hypervis_scaling = 0, laplace_scaling = 0;
  // hvcoord
  std::vector<double> hyam, hyai, hybm, hybi, dp0, nu_scale_top;
  double dvv[16], mp[16];
  // connectivity
  int max_corner = 1;
  std::vector<Connection> conns;
  // host staging of geometry (filled per element, uploaded in init_functors_c)
  std::vector<double> hD, hDinv, hmetdet, hrmetdet, hspheremp, hrspheremp, hfcor, hphis, hmetinv,
      htensorvisc, hvec, hgradphis;
  std::vector<int> hdss;
  // partition boundary (multi-GPU DSS): ghost element slots + per-peer entry lists
  int nghost = 0;
  std::vector<Peer> peers;
  std::vector<int> hsend_idx, hrecv_idx;       // local (elem*16+pt) to send / ghost (slot*16+pt) to fill
  int *d_send_idx = nullptr, *d_recv_idx = nullptr;
  double *d_sendbuf = nullptr, *d_recvbuf = nullptr;
  ncclComm_t comm = nullptr;
  ncclUniqueId nccl_id{};
  bool have_id = false;
  long halo_bytes = 0, exchanges = 0;
  // overlap of the exchange with the DSS-consuming kernel: elements without a ghost neighbour
  // (interior) are consumed while the halo is in flight on stream2, the others after it landed
  std::vector<int> h_interior, h_boundary;
  int *d_interior = nullptr, *d_boundary = nullptr;
  cudaStream_t stream2 = nullptr;
  cudaEvent_t ev_prod = nullptr, ev_halo = nullptr;
  // pipelined single-step service (thetal_b200_pipeline_step)
  StatePtrs pipe_in[2]{}, pipe_out[2]{};
  cudaStream_t s_h2d = nullptr, s_d2h = nullptr;
  cudaEvent_t ev_pin[4]{}, ev_pdone[4]{}, ev_pd2h[4]{};
  long pipe_calls = 0; bool pipe_ready = false;
  int ftype = -1;           // forcing algorithm (control_mod.F90): 0 and 2 apply the pushed forcing per model step
  double *d_ps_v[NTL] = {nullptr, nullptr, nullptr};   // elem%state%ps_v per time level, [ie][16]
  double *d_fm = nullptr, *d_fvtheta = nullptr, *d_fphi = nullptr;   // derived%FM (3 comps), FVTheta, FPHI
  bool have_forcing = false;
  int remap_alg = 1;        // vert_remap_u_alg (PPM variants only, as HOMMEXX)
  int do_remap = 1;         // prim_run_subcycle_c remaps after rsplit steps (thetal_b200_set_vertical_remap)
  int fuse_combo = 1;       // THETAL_B200_FUSE_COMBO=0: separate combo_kernel launch for the tstep_type 9 combination
  int fuse_dss = 1;         // THETAL_B200_FUSE_DSS=0: caar_dss_post_kernel between every CAAR stage and its DIRK solve
  int expl_cache = 1;       // THETAL_B200_EXPL_CACHE=0: the closing DIRK stage redoes the EOS of state(n0), state(nm1)
  int overlap = 1;          // halo of a DSS travels while the producer runs on the interior elements
  // device
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  std::vector<void *> allocs;
  StatePtrs tl[NTL]{}, tensA{}, tensB{};
  Geo geo{};
  Derived der{};
  RefStates ref{};
  int *d_err = nullptr, *d_maxiter = nullptr, *d_dirk_work = nullptr;
  long dirk_launches = 0;
  unsigned long long *d_maxdelta = nullptr, *d_extrema = nullptr;
  double *d_diag_part = nullptr;            // [nelem][NDIAG] element integrals
  unsigned long long *d_diag_acc = nullptr; // [NDIAG] max bits, [2] ps min/max bits, [NDIAG*3] limbs
  bool device_ready = false;
  long launches = 0;
  bool prof = false;
  double prof_ms[8] = {0};
  cudaEvent_t pe0 = nullptr, pe1 = nullptr;
  // DIRK statistics (element_state.F90:20-22)
  int max_itercnt = 0;
  double max_deltaerr = 0;
};

Session S;

void fail(int code, const std::string &msg)
{
  if (S.last_error == 0) {      // latched: the FIRST failure is what thetal_b200_last_error reports until it is cleared
    S.last_error = code;
    S.last_msg = msg;
  }
  fprintf(stderr, "thetal_b200: ERROR %d: %s\n", code, msg.c_str());
  if (S.abort_on_error) abort();
}

// Every check RETURNS from the entry point after fail() (which aborts unless abort_on_error == 0), so a
// latched error never lets the call run on with a broken precondition.  *_R variants for non-void functions.
#define CUDA_OK_R(x, ret) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fail(100, std::string(#x) + ": " + cudaGetErrorString(e_)); return ret; } } while (0)
#define CUDA_OK(x) CUDA_OK_R(x, )
#define REQUIRE_R(cond, what, ret) do { if (!(cond)) { fail(2, std::string("precondition failed: ") + (what)); return ret; } } while (0)
#define REQUIRE(cond, what) REQUIRE_R(cond, what, )

template <class T> T *dalloc(size_t n)
{
  void *p = nullptr;
  CUDA_OK_R(cudaMalloc(&p, n * sizeof(T)), nullptr);
  CUDA_OK_R(cudaMemset(p, 0, n * sizeof(T)), nullptr);
  S.allocs.push_back(p);
  return (T *)p;
}

StatePtrs alloc_state(int nelem, int nlev)
{
  StatePtrs s;
  s.v = dalloc<double>((size_t)nelem * 32 * nlev);
  s.w = dalloc<double>((size_t)nelem * 16 * (nlev + 1));
  s.th = dalloc<double>((size_t)nelem * 16 * nlev);
  s.phi = dalloc<double>((size_t)nelem * 16 * (nlev + 1));
  s.dp = dalloc<double>((size_t)nelem * 16 * nlev);
  return s;
}


// Errors::check_option equivalents (cxx_f90_interface_theta.cpp:57-73)
bool check_in(const char *name, int v, std::initializer_list<int> ok)
{
  for (int o : ok) if (o == v) return true;
  fail(3, std::string("init_simulation_params_c: unsupported value for ") + name + ": " + std::to_string(v));
  return false;
}
#define CHECK_IN(...) do { if (!check_in(__VA_ARGS__)) return; } while (0)
#define FAIL_IF(cond, code, msg) do { if (cond) { fail(code, msg); return; } } while (0)

}  // namespace

extern "C" {

// ---------------------------------------------------------------- session / comm
void initialize_hommexx_session(void)
{
  if (S.active) return;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) { fail(1, "no CUDA device: libthetal_b200 has no CPU fallback"); return; }
  CUDA_OK(cudaSetDevice(S.device % ndev));
  CUDA_OK(cudaStreamCreateWithFlags(&S.stream, cudaStreamNonBlocking));
  if (getenv("THETAL_B200_EXPL_CACHE")) S.expl_cache = atoi(getenv("THETAL_B200_EXPL_CACHE"));
  if (getenv("THETAL_B200_FUSE_COMBO")) S.fuse_combo = atoi(getenv("THETAL_B200_FUSE_COMBO"));
  if (getenv("THETAL_B200_FUSE_DSS")) S.fuse_dss = atoi(getenv("THETAL_B200_FUSE_DSS"));
  if (getenv("THETAL_B200_DIRK") && atoi(getenv("THETAL_B200_DIRK")) != 1) S.fuse_dss = 0;   // only the column kernel has the DSS form
  CUDA_OK(cudaEventCreate(&S.ev0));
  CUDA_OK(cudaEventCreate(&S.ev1));
  if (S.nranks > 1) {
    int prio_lo = 0, prio_hi = 0;
    CUDA_OK(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    CUDA_OK(cudaStreamCreateWithPriority(&S.stream2, cudaStreamNonBlocking, prio_hi));   // halo stream: its CTAs go first
    CUDA_OK(cudaEventCreateWithFlags(&S.ev_prod, cudaEventDisableTiming));
    CUDA_OK(cudaEventCreateWithFlags(&S.ev_halo, cudaEventDisableTiming));
    if (getenv("THETAL_B200_OVERLAP")) S.overlap = atoi(getenv("THETAL_B200_OVERLAP"));   // A/B switch, default 1
    if (!S.have_id) { fail(4, "nranks > 1 needs the ncclUniqueId of rank 0 (thetal_b200_init_comm)"); return; }
    ncclResult_t r = ncclCommInitRank(&S.comm, S.nranks, S.nccl_id, S.rank);
    if (r != ncclSuccess) { fail(5, std::string("ncclCommInitRank: ") + ncclGetErrorString(r)); return; }
  }
  S.active = true;
}

void finalize_hommexx_session(void)
{
  if (!S.active) return;
  cudaDeviceSynchronize();
  if (S.comm) { ncclCommDestroy(S.comm); S.comm = nullptr; }
  for (void *p : S.allocs) cudaFree(p);
  S.allocs.clear();
  if (S.ev0) cudaEventDestroy(S.ev0);
  if (S.ev1) cudaEventDestroy(S.ev1);
  if (S.stream) cudaStreamDestroy(S.stream);
  if (S.stream2) cudaStreamDestroy(S.stream2);
  if (S.ev_prod) cudaEventDestroy(S.ev_prod);
  if (S.ev_halo) cudaEventDestroy(S.ev_halo);
  if (S.pipe_ready) {
    cudaStreamDestroy(S.s_h2d); cudaStreamDestroy(S.s_d2h);
    for (int i = 0; i < 4; i++) { cudaEventDestroy(S.ev_pin[i]); cudaEventDestroy(S.ev_pdone[i]); cudaEventDestroy(S.ev_pd2h[i]); }
  }
  const int dev = S.device, rank = S.rank, nranks = S.nranks, aoe = S.abort_on_error;
  const bool have_id = S.have_id;
  const ncclUniqueId id = S.nccl_id;
  S = Session();
  S.device = dev; S.rank = rank; S.nranks = nranks; S.abort_on_error = aoe;
  S.have_id = have_id; S.nccl_id = id;
}

// mpi_cxx_f90_interface.cpp:20-23 converts the Fortran MPI communicator.  This library has no MPI: one rank
// drives one GPU and the ranks meet through NCCL, whose rank/size/id come from thetal_b200_init_comm (the
// Fortran shim passes par%rank, par%nprocs and the id rank 0 broadcasts over its own MPI_Bcast).  The handle is
// kept only so that a second call with a different communicator is reported.
void reset_cxx_comm(const int *f_comm)
{
  static int seen = -1;
  if (seen != -1 && f_comm && *f_comm != seen && S.active)
    fail(4, "reset_cxx_comm: communicator changed while a session is active; finalize_hommexx_session first");
  if (f_comm) seen = *f_comm;
}

void thetal_b200_init_comm(const int *rank, const int *nranks, const int *device, const void *nccl_id)
{
  S.rank = *rank; S.nranks = *nranks; S.device = *device;
  S.have_id = false;
  if (nccl_id) { memcpy(&S.nccl_id, nccl_id, sizeof(ncclUniqueId)); S.have_id = true; }
}
void thetal_b200_get_nccl_id(void *out)
{
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
  ncclUniqueId id;
  ncclResult_t r = ncclGetUniqueId(&id);
  if (r != ncclSuccess) { fail(5, std::string("ncclGetUniqueId: ") + ncclGetErrorString(r)); memset(out, 0, 128); return; }
  memcpy(out, &id, 128);
}
void thetal_b200_set_abort_on_error(const int *flag) { S.abort_on_error = *flag; }
int thetal_b200_last_error(void) { return S.last_error; }
const char *thetal_b200_last_error_msg(void) { return S.last_msg.c_str(); }
void thetal_b200_clear_error(void) { S.last_error = 0; S.last_msg.clear(); }
void thetal_b200_set_nlev(const int *nlev)
{
  if (*nlev < 2 || *nlev > 128) { fail(3, "nlev must be in [2,128]"); return; }
  S.nlev = *nlev;
}
void thetal_b200_set_extra_params(const int *unused, const int *nlev_tom)
{
  (void)unused;
  S.prm.nlev_tom = *nlev_tom;
}

// ---------------------------------------------------------------- connectivity
void init_connectivity(const int *num_local_elems, const int *max_corner_elems)
{
  S.conns.clear();
  // one neighbour per corner (uniform cubed sphere); refined meshes list several (dir_idx > 0 in the
  // reference's convert(), mpi_cxx_f90_interface.cpp:49-60) and would need a list per corner
  FAIL_IF(*max_corner_elems != 1, 3, "init_connectivity: max_corner_elems != 1 (regionally refined meshes) not supported");
  S.max_corner = *max_corner_elems;
  S.conns.reserve((size_t)*num_local_elems * 8);
}

void add_connection(const int *e1_lid, const int *e1_gid, const int *e1_pos, const int *e1_pid,
                    const int *e2_lid, const int *e2_gid, const int *e2_pos, const int *e2_pid)
{
  REQUIRE(*e1_lid >= 1 && *e1_gid >= 1 && *e1_pos >= 1 && *e1_pid >= 1 && *e2_lid >= 1 && *e2_gid >= 1 &&
          *e2_pos >= 1 && *e2_pid >= 1, "add_connection: F90 indices should start at 1");
  Connection c;
  c.e1 = *e1_lid - 1;
  c.e1_gid = *e1_gid - 1;
  c.dir1 = *e1_pos / S.max_corner - 1;      // W=0,E=1,S=2,N=3,SW=4,SE=5,NW=6,NE=7
  c.e2_lid = *e2_lid - 1; c.e2_gid = *e2_gid - 1; c.e2_pid = *e2_pid - 1;
  c.dir2 = *e2_pos / S.max_corner - 1;
  if (*e1_pid - 1 == S.rank) S.conns.push_back(c);
}

static void edge_points(int dir, int pts[4])
{ // natural ordering of the dofs of an edge (ConnectivityHelpers.hpp:30-47)
  for (int k = 0; k < 4; k++) {
    switch (dir) {
      case 0: pts[k] = 4 * k; break;          // W: i=0
      case 1: pts[k] = 4 * k + 3; break;      // E: i=3
      case 2: pts[k] = k; break;              // S: j=0
      default: pts[k] = 12 + k; break;        // N: j=3
    }
  }
}
static int corner_point(int dir) { return dir == 4 ? 0 : dir == 5 ? 3 : dir == 6 ? 12 : 15; }

void finalize_connectivity(void)
{
  const int nelem = S.prm.nelem;
  REQUIRE(nelem > 0, "init_elements_c must precede finalize_connectivity");
  // ghost slots for neighbours owned by other ranks, numbered in (rank, global id) order
  std::map<std::pair<int, int>, int> ghost;
  for (const Connection &c : S.conns)
    if (c.e2_pid != S.rank) {
      if (c.e2_pid < 0 || c.e2_pid >= S.nranks) { fail(4, "add_connection: neighbour rank outside the communicator"); return; }
      ghost[{c.e2_pid, c.e2_gid}] = 0;
    }
  int ng = 0;
  for (auto &g : ghost) g.second = nelem + ng++;
  S.nghost = ng;
  std::map<int, std::set<std::pair<int, int>>> sendset, recvset;   // per peer: (global id, point)
  std::map<int, int> lid_of_gid;
  // slot[e][p]: contributions by kind: 0..3 = W,E,S,N edge, 4 = corner
  std::vector<std::array<int, 5>> by_kind((size_t)nelem * 16);
  for (auto &a : by_kind) a.fill(-1);
  for (const Connection &c : S.conns) {
    const bool remote = c.e2_pid != S.rank;
    const int e2 = remote ? ghost[{c.e2_pid, c.e2_gid}] : c.e2_lid;
    if (remote) lid_of_gid[c.e1_gid] = c.e1;
    if (c.dir1 < 4) {
      REQUIRE(c.dir2 < 4, "edge connected to corner");
      int p1[4], p2[4];
      edge_points(c.dir1, p1); edge_points(c.dir2, p2);
      // orientation (ConnectivityHelpers.hpp:64-67): same-side pairs and W/N, E/S are reversed
      const int a = c.dir1, b = c.dir2;
      const bool rev = (a == b) || (a == 0 && b == 3) || (a == 3 && b == 0) || (a == 1 && b == 2) || (a == 2 && b == 1);
      for (int k = 0; k < 4; k++) {
        by_kind[(size_t)c.e1 * 16 + p1[k]][c.dir1] = e2 * 16 + p2[rev ? 3 - k : k];
        if (remote) { sendset[c.e2_pid].insert({c.e1_gid, p1[k]}); recvset[c.e2_pid].insert({c.e2_gid, p2[k]}); }
      }
    } else {
      by_kind[(size_t)c.e1 * 16 + corner_point(c.dir1)][4] = e2 * 16 + corner_point(c.dir2);
      if (remote) {
        sendset[c.e2_pid].insert({c.e1_gid, corner_point(c.dir1)});
        recvset[c.e2_pid].insert({c.e2_gid, corner_point(c.dir2)});
      }
    }
  }
  S.peers.clear(); S.hsend_idx.clear(); S.hrecv_idx.clear();
  for (auto &ps : sendset) {
    Peer pr;
    pr.rank = ps.first;
    pr.send_off = (int)S.hsend_idx.size();
    for (auto &gp : ps.second) S.hsend_idx.push_back(lid_of_gid[gp.first] * 16 + gp.second);
    pr.send_cnt = (int)ps.second.size();
    pr.recv_off = (int)S.hrecv_idx.size();
    for (auto &gp : recvset[pr.rank]) S.hrecv_idx.push_back(ghost[{pr.rank, gp.first}] * 16 + gp.second);
    pr.recv_cnt = (int)recvset[pr.rank].size();
    S.peers.push_back(pr);
  }
  // unpack order of edge_mod.F90:888-941: per i: S,N,W,E ; so SW: S,W  SE: E,S  NW: N,W  NE: N,E ; corners last
  S.hdss.assign((size_t)nelem * 48, -1);
  for (int e = 0; e < nelem; e++)
    for (int p = 0; p < 16; p++) {
      const auto &k = by_kind[(size_t)e * 16 + p];
      int order[5];
      if (p == 3) { order[0] = 1; order[1] = 2; order[2] = 0; order[3] = 3; }          // SE: E,S
      else { order[0] = 2; order[1] = 3; order[2] = 0; order[3] = 1; }                  // S,N,W,E
      order[4] = 4;
      int n = 0;
      for (int j = 0; j < 5; j++)
        if (k[order[j]] >= 0) { REQUIRE(n < 3, "more than 3 DSS neighbours at a point"); S.hdss[((size_t)e * 16 + p) * 3 + n++] = k[order[j]]; }
    }
  S.h_interior.clear(); S.h_boundary.clear();
  for (int e = 0; e < nelem; e++) {
    bool ghost_nbr = false;
    for (int j = 0; j < 48; j++) if (S.hdss[(size_t)e * 48 + j] >= nelem * 16) ghost_nbr = true;
    (ghost_nbr ? S.h_boundary : S.h_interior).push_back(e);
  }
}

}  // extern "C"

extern "C" {

// ---------------------------------------------------------------- parameters
void init_simulation_params_c(const int *remap_alg, const int *limiter_option, const int *rsplit,
    const int *qsplit, const int *time_step_type, const int *qsize, const int *state_frequency,
    const double *nu, const double *nu_p, const double *nu_q, const double *nu_s,
    const double *nu_div, const double *nu_top, const int *hypervis_order,
    const int *hypervis_subcycle, const int *hypervis_subcycle_tom,
    const double *hypervis_scaling, const double *laplace_scaling, const double *dcmip16_mu,
    const int *ftype, const int *theta_adv_form, const int *prescribed_wind,
    const int *use_moisture, const int *disable_diagnostics, const int *use_cpstar,
    const int *transport_alg, const int *theta_hydrostatic_mode, const char **test_case,
    const int *dt_remap_factor, const int *dt_tracer_factor, const double *scale_factor,
    const double *laplacian_rigid_factor, const int *nsplit, const int *pgrad_correction,
    const double *dp3d_thresh, const double *vtheta_thresh,
    const int *internal_diagnostics_level, const int *do_3d_turbulence,
    const double *tom_sponge_start)
{
  (void)limiter_option; (void)state_frequency; (void)nu_q; (void)use_moisture;
  (void)disable_diagnostics; (void)use_cpstar; (void)transport_alg; (void)test_case;
  (void)nsplit; (void)internal_diagnostics_level;
  (void)tom_sponge_start;
  CHECK_IN("remap_alg", *remap_alg, {1, 10, 11});                     // PPM only, like cxx_f90_interface_theta.cpp:57-73
  S.remap_alg = *remap_alg;
  CHECK_IN("hypervis_order", *hypervis_order, {2});
  CHECK_IN("time_step_type", *time_step_type, {1, 4, 5, 7, 9, 10});
  CHECK_IN("theta_advection_form", *theta_adv_form, {0, 1, 2});
  CHECK_IN("theta_hydrostatic_mode", *theta_hydrostatic_mode, {0});   // NH only (north star)
  CHECK_IN("prescribed_wind", *prescribed_wind, {0});
  CHECK_IN("qsize", *qsize, {0});                                     // tracers out of scope
  CHECK_IN("ftype", *ftype, {-1, 0, 2});
  CHECK_IN("do_3d_turbulence", *do_3d_turbulence, {0});
  FAIL_IF(!(*nu_p > 0) || !(*nu > 0) || !(*nu_div > 0), 3, "nu, nu_p, nu_div must be > 0");
  FAIL_IF(!(*dp3d_thresh > 0) || !(*vtheta_thresh > 0), 3, "dp3d_thresh, vtheta_thresh must be > 0");
  FAIL_IF(*hypervis_scaling < 0.0, 3, "hypervis_scaling must be >= 0");
  // tensor hyperviscosity needs the Cartesian vector laplacian, which has no nu_div (derivative_mod.F90:1146-1150)
  FAIL_IF(*hypervis_scaling > 0.0 && *nu_div != *nu, 3, "ERROR: tensorHV can not be used with nu_div/=nu");
  FAIL_IF(*laplace_scaling != 0.0, 3, "laplace_scaling != 0 not supported");
  FAIL_IF(*dcmip16_mu != 0.0, 3, "dcmip16_mu != 0 not supported");
  FAIL_IF(*rsplit < 0, 3, "rsplit must be >= 0 (0 = vertically Eulerian, PA:1227-1303)");
  FAIL_IF(*qsplit < 1, 3, "qsplit must be >= 1");
  FAIL_IF(fabs(*laplacian_rigid_factor * *scale_factor - 1.0) > 1e-12, 3,
          "planar geometry (laplacian_rigid_factor != 1/scale_factor) not supported");
  // dt_remap_factor / dt_tracer_factor (prim_driver.cpp:75-88): the dycore-only build has no tracers, so
  // the only consistent setting is dt_tracer_factor = qsplit, dt_remap_factor = rsplit*qsplit (what
  // namelist_mod.F90 derives when the namelist gives rsplit/qsplit); <= 0 means "derive it".
  // dt_remap_factor = 0 is the vertically Eulerian code (prim_driver.cpp:79-83), i.e. rsplit = 0.
  const int dtf = *dt_tracer_factor > 0 ? *dt_tracer_factor : *qsplit;
  const int drf = *dt_remap_factor > 0 ? *dt_remap_factor : *rsplit * dtf;
  FAIL_IF(drf != *rsplit * dtf, 3, "dt_remap_factor must equal rsplit*dt_tracer_factor (independent remap/tracer steps need tracers: out of scope)");
  S.ftype = *ftype;
  Params &p = S.prm;
  p.rsplit = *rsplit; p.qsplit = dtf; p.tstep_type = *time_step_type;
  p.theta_advect_form = *theta_adv_form; p.pgrad_correction = *pgrad_correction;
  p.hypervis_subcycle = *hypervis_subcycle; p.hypervis_subcycle_tom = *hypervis_subcycle_tom;
  p.hypervis_scaling_on = *hypervis_scaling > 0.0 ? 1 : 0;
  p.nu = *nu; p.nu_div = *nu_div; p.nu_p = *nu_p; p.nu_s = *nu_s; p.nu_top = *nu_top;
  p.dp3d_thresh = *dp3d_thresh; p.vtheta_thresh = *vtheta_thresh;
  p.rearth_inv = *laplacian_rigid_factor;     // == 1/scale_factor on the sphere (namelist_mod.F90:1045-1047)
  S.hypervis_order = *hypervis_order;
  S.hypervis_scaling = *hypervis_scaling; S.laplace_scaling = *laplace_scaling;
  S.params_set = true;
}

void init_elements_c(const int *nelemd)
{
  REQUIRE(S.nlev > 0, "thetal_b200_set_nlev must precede init_elements_c");
  const int n = *nelemd;
  S.prm.nelem = n; S.prm.nlev = S.nlev;
  S.hD.assign((size_t)n * 64, 0); S.hDinv.assign((size_t)n * 64, 0); S.hmetinv.assign((size_t)n * 64, 0);
  S.htensorvisc.assign((size_t)n * 64, 0); S.hvec.assign((size_t)n * 96, 0);
  S.hmetdet.assign((size_t)n * 16, 0); S.hrmetdet.assign((size_t)n * 16, 0); S.hspheremp.assign((size_t)n * 16, 0);
  S.hrspheremp.assign((size_t)n * 16, 0); S.hfcor.assign((size_t)n * 16, 0); S.hphis.assign((size_t)n * 16, 0);
  S.hgradphis.assign((size_t)n * 32, 0);
}

void init_hvcoord_c(const double *ps0, thetal_f90ptr hyam, thetal_f90ptr hyai, thetal_f90ptr hybm, thetal_f90ptr hybi)
{
  REQUIRE(S.nlev > 0, "thetal_b200_set_nlev must precede init_hvcoord_c");
  const int n = S.nlev;
  S.hyam.assign(*hyam, *hyam + n); S.hybm.assign(*hybm, *hybm + n);
  S.hyai.assign(*hyai, *hyai + n + 1); S.hybi.assign(*hybi, *hybi + n + 1);
  S.prm.ps0 = *ps0; S.prm.ptop = S.hyai[0] * *ps0;
  S.dp0.resize(n);
  for (int k = 0; k < n; k++)    // hybvcoord_mod.F90:184-187
    S.dp0[k] = (S.hyai[k + 1] - S.hyai[k]) * *ps0 + (S.hybi[k + 1] - S.hybi[k]) * *ps0;
}

void init_time_level_c(const int *nm1, const int *n0, const int *np1, const int *nstep, const int *nstep0)
{
  S.nm1 = *nm1 - 1; S.n0 = *n0 - 1; S.np1 = *np1 - 1; S.nstep = *nstep; S.nstep0 = *nstep0;
}

void init_reference_element_c(thetal_f90ptr deriv, thetal_f90ptr mass)
{
  memcpy(S.dvv, *deriv, sizeof S.dvv);
  memcpy(S.mp, *mass, sizeof S.mp);
}

void init_elements_2d_c(const int *ie, thetal_f90ptr D, thetal_f90ptr Dinv, thetal_f90ptr fcor,
    thetal_f90ptr spheremp, thetal_f90ptr rspheremp, thetal_f90ptr metdet, thetal_f90ptr metinv,
    thetal_f90ptr tensorvisc, thetal_f90ptr vec_sph2cart, const double *sphere_cart_vec,
    const double *sphere_latlon_vec, thetal_f90ptr tensorvisc2)
{
  (void)sphere_cart_vec; (void)sphere_latlon_vec; (void)tensorvisc2;
  const size_t e = *ie;
  REQUIRE((int)e < S.prm.nelem, "init_elements_2d_c: ie out of range");
  memcpy(&S.hD[e * 64], *D, 64 * sizeof(double));
  memcpy(&S.hDinv[e * 64], *Dinv, 64 * sizeof(double));
  memcpy(&S.hmetinv[e * 64], *metinv, 64 * sizeof(double));
  if (tensorvisc && *tensorvisc) memcpy(&S.htensorvisc[e * 64], *tensorvisc, 64 * sizeof(double));
  if (vec_sph2cart && *vec_sph2cart) memcpy(&S.hvec[e * 96], *vec_sph2cart, 96 * sizeof(double));
  memcpy(&S.hfcor[e * 16], *fcor, 16 * sizeof(double));
  memcpy(&S.hspheremp[e * 16], *spheremp, 16 * sizeof(double));
  memcpy(&S.hrspheremp[e * 16], *rspheremp, 16 * sizeof(double));
  memcpy(&S.hmetdet[e * 16], *metdet, 16 * sizeof(double));
  for (int p = 0; p < 16; p++) S.hrmetdet[e * 16 + p] = 1.0 / S.hmetdet[e * 16 + p];   // cube_mod.F90:274
}

void init_tensorvisc_c(const int *ie, thetal_f90ptr tensorvisc)
{
  memcpy(&S.htensorvisc[(size_t)*ie * 64], *tensorvisc, 64 * sizeof(double));
}
void init_tensorvisc2_c(const int *ie, thetal_f90ptr tensorvisc2) { (void)ie; (void)tensorvisc2; }

void init_geopotential_c(const int *ie, thetal_f90ptr phis, thetal_f90ptr gradphis)
{
  const size_t e = *ie;
  memcpy(&S.hphis[e * 16], *phis, 16 * sizeof(double));
  memcpy(&S.hgradphis[e * 32], *gradphis, 32 * sizeof(double));
}

void init_diagnostics_c(thetal_f90ptr, thetal_f90ptr, thetal_f90ptr, thetal_f90ptr, thetal_f90ptr,
                        thetal_f90ptr, thetal_f90ptr) {}

// Create device objects: geometry upload, state/tens allocation (cxx_f90_interface_theta.cpp:568-640)
void init_functors_c(const int *allocate_buffer)
{
  (void)allocate_buffer;
  if (!S.active) initialize_hommexx_session();
  if (!S.active) return;
  REQUIRE(S.params_set, "init_simulation_params_c must precede init_functors_c");
  REQUIRE(!S.hyai.empty(), "init_hvcoord_c must precede init_functors_c");
  REQUIRE(!S.hdss.empty(), "finalize_connectivity must precede init_functors_c");
  if (S.device_ready) return;
  const int n = S.prm.nelem, nlev = S.nlev;
  auto up = [&](const std::vector<double> &h) -> const double * {
    double *d = dalloc<double>(h.size());
    if (d) CUDA_OK_R(cudaMemcpy(d, h.data(), h.size() * sizeof(double), cudaMemcpyHostToDevice), nullptr);
    return d;
  };
  S.geo.D = up(S.hD); S.geo.Dinv = up(S.hDinv); S.geo.metinv = up(S.hmetinv);
  S.geo.tensorvisc = up(S.htensorvisc); S.geo.vec_sph2cart = up(S.hvec);
  S.geo.metdet = up(S.hmetdet); S.geo.rmetdet = up(S.hrmetdet); S.geo.spheremp = up(S.hspheremp);
  S.geo.rspheremp = up(S.hrspheremp); S.geo.fcor = up(S.hfcor); S.geo.phis = up(S.hphis);
  S.geo.gradphis = up(S.hgradphis);
  int *dd = dalloc<int>(S.hdss.size());
  CUDA_OK(cudaMemcpy(dd, S.hdss.data(), S.hdss.size() * sizeof(int), cudaMemcpyHostToDevice));
  S.geo.dss_src = dd;
  for (int t = 0; t < NTL; t++) S.tl[t] = alloc_state(n, nlev);
  S.tensA = alloc_state(n + S.nghost, nlev);      // ghost element slots behind the local elements
  S.tensB = alloc_state(n + S.nghost, nlev);
  if (!S.peers.empty()) {
    const size_t nlyr = 6 * (size_t)nlev + 1;
    S.d_send_idx = dalloc<int>(S.hsend_idx.size());
    S.d_recv_idx = dalloc<int>(S.hrecv_idx.size());
    CUDA_OK(cudaMemcpy(S.d_send_idx, S.hsend_idx.data(), S.hsend_idx.size() * sizeof(int), cudaMemcpyHostToDevice));
    CUDA_OK(cudaMemcpy(S.d_recv_idx, S.hrecv_idx.data(), S.hrecv_idx.size() * sizeof(int), cudaMemcpyHostToDevice));
    S.d_sendbuf = dalloc<double>(S.hsend_idx.size() * nlyr);
    S.d_recvbuf = dalloc<double>(S.hrecv_idx.size() * nlyr);
    S.d_interior = dalloc<int>(S.h_interior.size() + 1);
    S.d_boundary = dalloc<int>(S.h_boundary.size() + 1);
    CUDA_OK(cudaMemcpy(S.d_interior, S.h_interior.data(), S.h_interior.size() * sizeof(int), cudaMemcpyHostToDevice));
    CUDA_OK(cudaMemcpy(S.d_boundary, S.h_boundary.data(), S.h_boundary.size() * sizeof(int), cudaMemcpyHostToDevice));
  }
  S.der.vn0 = dalloc<double>((size_t)n * 32 * nlev);
  S.der.eta_dot_dpdn = dalloc<double>((size_t)n * 16 * (nlev + 1));
  S.der.omega_p = dalloc<double>((size_t)n * 16 * nlev);
  S.ref.theta_ref = dalloc<double>((size_t)n * 16 * nlev);
  S.ref.dp_ref = dalloc<double>((size_t)n * 16 * nlev);
  S.ref.phi_ref = dalloc<double>((size_t)n * 16 * (nlev + 1));
  for (int t = 0; t < NTL; t++) S.d_ps_v[t] = dalloc<double>((size_t)n * 16);
  S.d_dirk_work = dalloc<int>(2);
  S.d_err = dalloc<int>(1); S.d_maxiter = dalloc<int>(1); S.d_maxdelta = dalloc<unsigned long long>(1);
  S.d_diag_part = dalloc<double>((size_t)S.prm.nelem * NDIAG); S.d_diag_acc = dalloc<unsigned long long>(NDIAG + 2 + NDIAG * thetal_reprosum::NLIMB);
  if (S.nu_scale_top.empty()) S.nu_scale_top.assign(nlev, 0.0);
  upload_constants(S.dvv, S.mp, S.hybi.data(), S.dp0.data(), S.nu_scale_top.data(), nlev);
  upload_hyai(S.hyai.data(), nlev);
  CUDA_OK(cudaDeviceSynchronize());
  S.device_ready = true;
}

// The exchange objects (gather table, per-peer halo lists and buffers) are created by finalize_connectivity and
// init_functors_c; nothing is left to do here (theta_f2c_mod.F90:176-178 calls it after init_functors_c).
void init_boundary_exchanges_c(void)
{
  REQUIRE(S.device_ready, "init_functors_c must precede init_boundary_exchanges_c");
}
// prim_driver.cpp:31-57: dp3d(n0) from ps_v(n0) on the reference levels
void initialize_dp3d_from_ps_c(void)
{
  REQUIRE(S.device_ready && S.n0 >= 0 && S.n0 < NTL, "initialize_dp3d_from_ps_c before init_functors_c / init_time_level_c");
  launch_dp_from_ps(S.tl[S.n0].dp, S.d_ps_v[S.n0], S.prm.nelem, S.nlev, S.prm.ps0, S.stream);
  S.launches += 1;
  thetal_b200_sync();
}

}  // extern "C"

namespace {

void copy_state_h2d(const double *h, double *const d[NTL], size_t per_elem)
{ // h: [ie][tl][per_elem]  ->  d[tl]: [ie][per_elem]
  const int n = S.prm.nelem;
  for (int t = 0; t < NTL; t++)
    CUDA_OK(cudaMemcpy2DAsync(d[t], per_elem * sizeof(double), h + t * per_elem, NTL * per_elem * sizeof(double),
                              per_elem * sizeof(double), n, cudaMemcpyHostToDevice, S.stream));
}
void copy_state_d2h(double *h, double *const d[NTL], size_t per_elem)
{
  const int n = S.prm.nelem;
  for (int t = 0; t < NTL; t++)
    CUDA_OK(cudaMemcpy2DAsync(h + t * per_elem, NTL * per_elem * sizeof(double), d[t], per_elem * sizeof(double),
                              per_elem * sizeof(double), n, cudaMemcpyDeviceToHost, S.stream));
}

void check_device_errors(const char *where)
{
  int err = 0;
  CUDA_OK(cudaMemcpyAsync(&err, S.d_err, sizeof(int), cudaMemcpyDeviceToHost, S.stream));
  CUDA_OK(cudaStreamSynchronize(S.stream));
  if (err & ERR_EOS_BAD_STATE) {     // eos.F90:128-145 aborts the job
    CUDA_OK(cudaMemsetAsync(S.d_err, 0, sizeof(int), S.stream));
    fail(10, std::string("EOS bad state: d(phi), dp3d or vtheta_dp < 0 (detected in ") + where + ")");
  }
  if (err & ERR_REMAP_NEGATIVE_LAYER) {   // vertremap_mod.F90:87-103 aborts the job
    CUDA_OK(cudaMemsetAsync(S.d_err, 0, sizeof(int), S.stream));
    fail(11, std::string("negative layer thickness.  timestep or remap time too large (detected in ") + where + ")");
  }
  if (err & ERR_DIRK_NOT_CONVERGED) {  // imex_mod.F90:353-357: warning only
    fprintf(stderr, "thetal_b200: WARNING: IMEX solver failed b/c max iteration count was met\n");
    CUDA_OK(cudaMemsetAsync(S.d_err, 0, sizeof(int), S.stream));
  }
}

struct ProfScope {   // per-family device time when thetal_b200_profile_step is active
  int fam;
  explicit ProfScope(int f) : fam(f) { if (S.prof) cudaEventRecord(S.pe0, S.stream); }
  ~ProfScope() {
    if (!S.prof) return;
    cudaEventRecord(S.pe1, S.stream); cudaEventSynchronize(S.pe1);
    float ms = 0; cudaEventElapsedTime(&ms, S.pe0, S.pe1); S.prof_ms[fam] += ms;
  }
};

struct Nvtx {        // t_startf/t_stopf of the reference (prim_advance_mod.F90:116,591,670,1105,1720; imex_mod.F90:140;
  explicit Nvtx(const char *name) { nvtxRangePushA(name); }     // viscosity_theta.F90:130), visible to nsys/ncu
  ~Nvtx() { nvtxRangePop(); }
};

// ---- bndry_exchangeV (bndry_mod.F90:51-169) for one DSS: pack the partition-boundary points of
// `t`, ncclSend/ncclRecv with every neighbour rank, unpack into the ghost slots of `t`.
// nk levels of each field (+ w_i(nlevp) when wbot); all on S.stream, so it is ordered between the
// kernel that produced `t` and the DSS kernel that gathers from it.
#define NCCL_OK(x) do { ncclResult_t r_ = (x); if (r_ != ncclSuccess) { fail(5, std::string(#x) + ": " + ncclGetErrorString(r_)); return; } } while (0)
void halo_exchange(const StatePtrs &t, int nk, int wbot, cudaStream_t st)
{
  if (S.peers.empty()) return;
  const int nlev = S.nlev, nlyr = 6 * nk + wbot;
  launch_halo_pack(t, S.d_send_idx, (int)S.hsend_idx.size(), nk, wbot, nlev, S.d_sendbuf, st);
  NCCL_OK(ncclGroupStart());
  for (const Peer &p : S.peers) {
    NCCL_OK(ncclSend(S.d_sendbuf + (size_t)p.send_off * nlyr, (size_t)p.send_cnt * nlyr, ncclDouble, p.rank, S.comm, st));
    NCCL_OK(ncclRecv(S.d_recvbuf + (size_t)p.recv_off * nlyr, (size_t)p.recv_cnt * nlyr, ncclDouble, p.rank, S.comm, st));
  }
  NCCL_OK(ncclGroupEnd());
  launch_halo_unpack(t, S.d_recv_idx, (int)S.hrecv_idx.size(), nk, wbot, nlev, S.d_recvbuf, st);
  S.launches += 2;
  S.exchanges += 1;
  S.halo_bytes += (long)(S.hsend_idx.size() + S.hrecv_idx.size()) * nlyr * 8;
}

// One DSS across ranks.  `produce(elist, count)` launches, on the stream it is given, the kernel that writes
// the pre-DSS buffer `t` for the listed elements (null = all).  The partition-boundary elements are produced
// first, on the high-priority halo stream, followed there by the pack and the ncclSend/ncclRecv group; the
// interior elements are produced meanwhile on the main stream, which then waits for the halo.  The kernel
// that gathers from `t` is launched by the caller on the main stream for all elements.  (Reading the
// contributions of other ranks straight from the receive buffer instead of ghost slots was measured and
// dropped: the second address form in dss_sum stops the compiler from batching the neighbour loads and
// doubles the time of every gather kernel, 0.80 -> 1.75 ms per step for the five DSS/post launches.)
template <class F> void produce_and_exchange(const StatePtrs &t, int nk, int wbot, F produce)
{
  if (S.peers.empty()) { produce(nullptr, 0, S.stream); return; }
  if (!S.overlap) { produce(nullptr, 0, S.stream); halo_exchange(t, nk, wbot, S.stream); return; }
  CUDA_OK(cudaEventRecord(S.ev_prod, S.stream));
  CUDA_OK(cudaStreamWaitEvent(S.stream2, S.ev_prod, 0));
  produce(S.d_boundary, (int)S.h_boundary.size(), S.stream2);
  halo_exchange(t, nk, wbot, S.stream2);
  CUDA_OK(cudaEventRecord(S.ev_halo, S.stream2));
  produce(S.d_interior, (int)S.h_interior.size(), S.stream);
  CUDA_OK(cudaStreamWaitEvent(S.stream, S.ev_halo, 0));
  S.launches += 1;
}

// ---- the pieces of prim_advance_exp, enqueued on S.stream --------------------------------
// save_expl: 0 = no, 1 / 2 = keep the explicit stiff terms of state(n0) for the closing DIRK stage of this step in the
// (otherwise idle) hyperviscosity scratch tensB: w/phi slots (1) or dp/th slots (2)
void run_caar(int np1, int nm1, int n0, double dt2, double eta_ave_w, double s1, double s2, double s3, int save_expl = 0,
              int combo_b = -1, int combo_c = -1, bool defer_dss = false)
{
  Nvtx r("compute_andor_apply_rhs");
  CaarArgs a;
  if (save_expl == 1) { a.ew = S.tensB.w; a.ephi = S.tensB.phi; }
  if (save_expl == 2) { a.ew = S.tensB.dp; a.ephi = S.tensB.th; }
  a.n0 = S.tl[n0]; a.nm1 = S.tl[nm1]; a.tens = S.tensA; a.geo = S.geo; a.der = S.der; a.prm = S.prm;
  a.dt2 = dt2; a.eta_ave_w = eta_ave_w; a.scale1 = s1; a.scale2 = s2; a.scale3 = s3; a.err = S.d_err;
  {
    ProfScope ps(0);
    Nvtx rx("caar_bexchV");
    produce_and_exchange(S.tensA, S.nlev, 1, [&](const int *el, int n, cudaStream_t st) {
      a.elist = el; a.ecount = n; launch_caar_pre(a, st); });
  }
  if (defer_dss) { S.launches += 1; return; }      // the DIRK launch that follows gathers from tensA itself
  DssArgs d;
  d.tens = S.tensA; d.out = S.tl[np1]; d.geo = S.geo; d.prm = S.prm; d.dt2 = dt2; d.scale1 = s1; d.scale3 = s3;
  d.do_limiter = (s3 != 0.0);
  d.elist = nullptr; d.ecount = 0;
  if (combo_b >= 0) { d.combo_b = S.tl[combo_b]; d.combo_c = S.tl[combo_c]; }    // PA:313-326 fused behind the DSS
  { ProfScope ps(1); launch_caar_dss_post(d, S.stream); }
  S.launches += 2;
}

// cached: the explicit terms of state(n0) were saved by run_caar(..., save_expl = 1), those of state(nm1) by save_expl = 2
// dss_dt2 > 0: the preceding run_caar(..., defer_dss = true) left its pre-DSS result in tensA; its dt2, scale1 and scale3
void run_dirk(int nm1, double a_nm1, int n0, double a_n0, int np1, double dt2, bool cached = false,
              double dss_dt2 = 0.0, double dss_s1 = 0.0, double dss_s3 = 0.0)
{
  Nvtx r("compute_stage_value_dirk");
  DirkArgs a;
  if (dss_dt2 > 0.0) { a.tens = S.tensA; a.dss_dt2 = dss_dt2; a.dss_scale1 = dss_s1; a.dss_limiter = (dss_s3 != 0.0); }
  if (cached && S.expl_cache && S.prm.rsplit > 0) { a.ew_n0 = S.tensB.w; a.ephi_n0 = S.tensB.phi; a.ew_nm1 = S.tensB.dp; a.ephi_nm1 = S.tensB.th; }
  a.np1 = S.tl[np1]; a.n0 = S.tl[n0]; a.nm1 = S.tl[nm1]; a.geo = S.geo; a.prm = S.prm;
  a.dt2 = dt2; a.alphadt_n0 = a_n0; a.alphadt_nm1 = a_nm1;
  a.err = S.d_err; a.max_iter = S.d_maxiter; a.max_deltaerr_bits = S.d_maxdelta;
  a.work = S.d_dirk_work + (S.dirk_launches & 1); a.work_other = S.d_dirk_work + ((S.dirk_launches + 1) & 1);
  S.dirk_launches++;
  { ProfScope ps(2); launch_dirk(a, S.stream); }
  S.launches += 1;
}

// One IMEX stage whose DIRK solve works on the time level the CAAR stage has just produced (PA:283-311, 339-361):
// compute_andor_apply_rhs(np1, nm1, n0, dt2, ...) then compute_stage_value_dirk(d_nm1, a_nm1, d_n0, a_n0, np1, d_dt2).
// With fuse_dss the DSS between the two is done by the DIRK kernel (dirk_col_kernel<.., DSS = true>).
void run_stage(int np1, int nm1, int n0, double dt2, double eta_ave_w, double s1, double s2, double s3, int save_expl,
               int d_nm1, double a_nm1, int d_n0, double a_n0, double d_dt2)
{
  if (S.fuse_dss && dt2 > 0.0) {
    run_caar(np1, nm1, n0, dt2, eta_ave_w, s1, s2, s3, save_expl, -1, -1, true);
    run_dirk(d_nm1, a_nm1, d_n0, a_n0, np1, d_dt2, false, dt2, s1, s3);
  } else {
    run_caar(np1, nm1, n0, dt2, eta_ave_w, s1, s2, s3, save_expl);
    run_dirk(d_nm1, a_nm1, d_n0, a_n0, np1, d_dt2);
  }
}

void run_hypervis(int nt, double dt2, double eta_ave_w)
{
  ProfScope ps(3);
  Nvtx r("advance_hypervis");
  (void)eta_ave_w;   // dpdiss_ave / dpdiss_biharmonic diagnostics (PA:651-656) are not carried
  HvArgs a;
  a.st = S.tl[nt]; a.geo = S.geo; a.ref = S.ref; a.prm = S.prm;
  const int hvs = S.prm.hypervis_subcycle;
  // viscosity_theta.F90:77-83: nu_ratio1 = nu_div/nu on the first laplacian, nu_ratio2 = 1
  const double nu_ratio1 = (S.prm.nu_div != S.prm.nu) ? S.prm.nu_div / S.prm.nu : 1.0;
  for (int ic = 0; ic < hvs; ic++) {
    a.dt = dt2 / hvs;
    a.th_is_theta = ic > 0; a.last_subcycle = (ic == hvs - 1);
    a.nu_ratio = nu_ratio1; a.tin = S.tensB; a.tout = S.tensA;
    { Nvtx rx("biwkdp3d_bexchV");
      produce_and_exchange(S.tensA, S.nlev, 0, [&](const int *el, int n, cudaStream_t st) { a.elist = el; a.ecount = n; launch_hv(0, a, st); }); }
    a.nu_ratio = 1.0;       a.tin = S.tensA; a.tout = S.tensB;
    { Nvtx rx("ahdp_bexchV2");
      produce_and_exchange(S.tensB, S.nlev, 0, [&](const int *el, int n, cudaStream_t st) { a.elist = el; a.ecount = n; launch_hv(1, a, st); }); }
    a.tin = S.tensB; a.tout = S.tensA;
    a.elist = nullptr; a.ecount = 0;
    launch_hv(2, a, S.stream);
    S.launches += 3;
  }
  if (hvs == 0) fail(3, "hypervis_subcycle must be >= 1");
  if (S.prm.nu_top > 0 && S.prm.hypervis_subcycle_tom > 0 && S.prm.nlev_tom > 0) {   // PA:780
    const int nt_ = S.prm.hypervis_subcycle_tom;
    for (int ic = 0; ic < nt_; ic++) {
      a.dt = dt2 / nt_; a.th_is_theta = 0; a.last_subcycle = 0;
      a.tin = S.tensB; a.tout = S.tensA;
      produce_and_exchange(S.tensA, S.prm.nlev_tom, 0, [&](const int *el, int n, cudaStream_t st) { a.elist = el; a.ecount = n; launch_hv(3, a, st); });
      a.tin = S.tensA; a.tout = S.tensB;
      a.elist = nullptr; a.ecount = 0;
      launch_hv(4, a, S.stream);
      S.launches += 2;
    }
  }
}

void run_combo(int kind, int ia, int ib, int ic)
{
  ComboArgs c;
  c.a = S.tl[ia]; c.b = S.tl[ib]; c.c = S.tl[ic]; c.geo = S.geo; c.prm = S.prm; c.kind = kind;
  { ProfScope ps(4); launch_combo(c, S.stream); }
  S.launches += 1;
}

// PA:81-513
void run_prim_advance_exp(double dt, int nm1, int n0, int np1)
{
  Nvtx r("prim_advance_exp");
  const double eta_ave_w = 1.0 / S.prm.qsplit;                                       // PA:137
  launch_wbc(S.tl[n0], S.geo, S.prm.nelem, S.prm.nlev, S.stream);                    // PA:140-143
  S.launches += 1;
  switch (S.prm.tstep_type) {
    case 1:
      run_caar(np1, n0, n0, dt / 2, 0.0, 1, 1, 1);
      run_caar(np1, n0, np1, dt, eta_ave_w, 1, 1, 1);
      break;
    case 4:
      run_caar(np1, n0, n0, dt / 4, 0.0, 1, 1, 1);
      run_caar(np1, n0, np1, dt / 6, 0.0, 1, 1, 1);
      run_caar(np1, n0, np1, 3 * dt / 8, 0.0, 1, 1, 1);
      run_caar(np1, n0, np1, dt / 2, 0.0, 1, 1, 1);
      run_caar(np1, n0, np1, dt, eta_ave_w * 1.0, 1, 1, 1);
      break;
    case 5:
      run_caar(nm1, n0, n0, dt / 5, eta_ave_w / 4, 1, 1, 1);
      run_caar(np1, n0, nm1, dt / 5, 0.0, 1, 1, 1);
      run_caar(np1, n0, np1, dt / 3, 0.0, 1, 1, 1);
      run_caar(np1, n0, np1, 2 * dt / 3, 0.0, 1, 1, 1);
      run_combo(5, nm1, n0, n0);
      run_caar(np1, nm1, np1, 3 * dt / 4, 3 * eta_ave_w / 4, 1, 1, 1);
      break;
    case 7: {
      const double a1 = 0.0, a2 = 1 - a1;
      double dt2 = dt / 4;
      run_stage(np1, n0, n0, dt2, 0.0, 1, 0, 1, 0,  nm1, 0.0, n0, a1 * dt2, a2 * dt2);
      dt2 = dt / 6;
      run_stage(nm1, n0, np1, dt2, 0.0, 1, 0, 1, 0, nm1, 0.0, n0, a1 * dt2, a2 * dt2);
      dt2 = 3 * dt / 8;
      run_stage(np1, n0, nm1, dt2, 0.0, 1, 0, 1, 0, nm1, 0.0, n0, a1 * dt2, a2 * dt2);
      dt2 = dt / 2;
      run_stage(np1, n0, np1, dt2, 0.0, 1, 0, 1, 0, nm1, 0.0, n0, a1 * dt2, a2 * dt2);
      run_stage(np1, n0, np1, dt, eta_ave_w * 1.0, 1, 0, 1, 0, nm1, 0.0, n0, a1 * dt, a2 * dt);
      break;
    }
    case 9: {
      run_stage(nm1, n0, n0, dt / 5, eta_ave_w / 4, 1, 0, 1, 1,  nm1, 0.0, n0, 0.0, dt / 5);
      run_stage(np1, n0, nm1, dt / 5, 0.0, 1, 0, 1, 2,           nm1, 0.0, n0, 0.0, dt / 5);
      run_stage(np1, n0, np1, dt / 3, 0.0, 1, 0, 1, 0,           nm1, 0.0, n0, 0.0, dt / 3);
      run_stage(np1, n0, np1, 2 * dt / 3, 0.0, 1, 0, 1, 0,       nm1, 0.0, n0, 0.0, 2 * dt / 3);
      if (S.fuse_combo) run_caar(np1, nm1, np1, 3 * dt / 4, 3 * eta_ave_w / 4, 1, 0, 1, 0, nm1, n0);
      else { run_caar(np1, nm1, np1, 3 * dt / 4, 3 * eta_ave_w / 4, 1, 0, 1); run_combo(9, np1, nm1, n0); }
      const double a1 = 5 * dt / 18, a2 = dt / 36, a3 = 8 * dt / 18;
      run_dirk(nm1, a2, n0, a1, np1, a3, true);     // neither state(n0) nor state(nm1) changed since stages 1 and 2
      break;
    }
    case 10: {
      double dt2 = dt / 4;
      run_stage(nm1, n0, n0, dt2, 0.0, 1, 0, 1, 1, nm1, 0.0, n0, 0.0, dt2);
      dt2 = dt / 6;
      run_stage(np1, n0, nm1, dt2, 0.0, 1, 0, 1, 2, nm1, 0.0, n0, 0.0, dt2);
      dt2 = 3 * dt / 8;
      run_stage(np1, n0, np1, dt2, 0.0, 1, 0, 1, 0, nm1, 0.0, n0, 0.0, dt2);
      dt2 = dt / 2;
      run_stage(np1, n0, np1, dt2, 0.0, 1, 0, 1, 0, nm1, 0.0, n0, 0.0, dt2);
      run_caar(np1, n0, np1, dt, eta_ave_w * 1.0, 1, 0, 1);
      const double a1 = .24362, a2 = .34184, a3 = 1 - (a1 + a2);
      run_dirk(nm1, a2 * dt, n0, a1 * dt, np1, a3 * dt, true);
      break;
    }
    default: fail(3, "bad choice of tstep_type");
  }
  if (S.hypervis_order == 2 && S.prm.nu > 0) run_hypervis(np1, dt, eta_ave_w);          // PA:499-500
}

}  // namespace

extern "C" {

// ---------------------------------------------------------------- state transfer
void init_elements_states_c(thetal_f90ptr v, thetal_f90ptr w_i, thetal_f90ptr vtheta_dp,
    thetal_f90ptr phinh_i, thetal_f90ptr dp3d, thetal_f90ptr qdp, thetal_f90ptr ps_v)
{
  (void)qdp;
  REQUIRE(S.device_ready, "init_functors_c must precede init_elements_states_c");
  if (!S.device_ready) return;
  const size_t nlev = S.nlev;
  double *dv[NTL], *dw[NTL], *dth[NTL], *dphi[NTL], *ddp[NTL];
  for (int t = 0; t < NTL; t++) { dv[t] = S.tl[t].v; dw[t] = S.tl[t].w; dth[t] = S.tl[t].th; dphi[t] = S.tl[t].phi; ddp[t] = S.tl[t].dp; }
  copy_state_h2d(*v, dv, 32 * nlev);
  copy_state_h2d(*w_i, dw, 16 * (nlev + 1));
  copy_state_h2d(*vtheta_dp, dth, 16 * nlev);
  copy_state_h2d(*phinh_i, dphi, 16 * (nlev + 1));
  copy_state_h2d(*dp3d, ddp, 16 * nlev);
  if (ps_v && *ps_v) copy_state_h2d(*ps_v, S.d_ps_v, 16);          // ps_v(np,np,timelevels,nelemd)
  CUDA_OK(cudaStreamSynchronize(S.stream));
}

// Addition: the PLAIN theta-l build keeps the state inside an array of element structs
// (element_mod.F90:32-144, elem(ie)%state%v(np,np,2,nlev,timelevels) ...).  base pointers = the field of
// elem(1), elem_stride_bytes = storage size of one element_t: no Fortran-side copy into flat arrays.
void init_elements_states_strided_c(thetal_f90ptr v, thetal_f90ptr w_i, thetal_f90ptr vtheta_dp,
    thetal_f90ptr phinh_i, thetal_f90ptr dp3d, const long *elem_stride_bytes)
{
  REQUIRE(S.device_ready, "init_functors_c must precede init_elements_states_strided_c");
  const size_t nlev = S.nlev, n = S.prm.nelem, stride = (size_t)*elem_stride_bytes;
  REQUIRE(stride >= NTL * 32 * nlev * sizeof(double), "init_elements_states_strided_c: stride smaller than one element's v");
  auto copy = [&](const double *h, int which, size_t per_elem) {
    for (int t = 0; t < NTL; t++) {
      double *d = which == 0 ? S.tl[t].v : which == 1 ? S.tl[t].w : which == 2 ? S.tl[t].th : which == 3 ? S.tl[t].phi : S.tl[t].dp;
      CUDA_OK(cudaMemcpy2DAsync(d, per_elem * sizeof(double), h + t * per_elem, stride, per_elem * sizeof(double), n,
                                cudaMemcpyHostToDevice, S.stream));
    }
  };
  copy(*v, 0, 32 * nlev); copy(*w_i, 1, 16 * (nlev + 1)); copy(*vtheta_dp, 2, 16 * nlev);
  copy(*phinh_i, 3, 16 * (nlev + 1)); copy(*dp3d, 4, 16 * nlev);
  CUDA_OK(cudaStreamSynchronize(S.stream));
}

void cxx_push_results_to_f90(thetal_f90ptr_out v, thetal_f90ptr_out w_i, thetal_f90ptr_out vtheta_dp,
    thetal_f90ptr_out phinh_i, thetal_f90ptr_out dp3d, thetal_f90ptr_out ps_v, thetal_f90ptr_out qdp,
    thetal_f90ptr_out q, thetal_f90ptr_out omega_p)
{
  (void)qdp; (void)q;
  REQUIRE(S.device_ready, "no device state");
  if (!S.device_ready) return;
  const size_t nlev = S.nlev;
  double *dv[NTL], *dw[NTL], *dth[NTL], *dphi[NTL], *ddp[NTL];
  for (int t = 0; t < NTL; t++) { dv[t] = S.tl[t].v; dw[t] = S.tl[t].w; dth[t] = S.tl[t].th; dphi[t] = S.tl[t].phi; ddp[t] = S.tl[t].dp; }
  copy_state_d2h(*v, dv, 32 * nlev);
  copy_state_d2h(*w_i, dw, 16 * (nlev + 1));
  copy_state_d2h(*vtheta_dp, dth, 16 * nlev);
  copy_state_d2h(*phinh_i, dphi, 16 * (nlev + 1));
  copy_state_d2h(*dp3d, ddp, 16 * nlev);
  if (ps_v && *ps_v) {         // ps_v = hyai(1) ps0 + sum dp3d of every time level (vertremap_mod.F90:79-84)
    for (int t = 0; t < NTL; t++) launch_ps_from_dp(S.tl[t].dp, S.d_ps_v[t], S.prm.nelem, S.nlev, S.prm.ptop, S.stream);
    S.launches += NTL;
    copy_state_d2h(*ps_v, S.d_ps_v, 16);
  }
  if (omega_p && *omega_p)
    CUDA_OK(cudaMemcpyAsync(*omega_p, S.der.omega_p, (size_t)S.prm.nelem * 16 * nlev * sizeof(double), cudaMemcpyDeviceToHost, S.stream));
  CUDA_OK(cudaStreamSynchronize(S.stream));
  check_device_errors("cxx_push_results_to_f90");
}

void push_test_state_to_c(thetal_f90ptr ps_v, thetal_f90ptr dp3d, thetal_f90ptr vtheta_dp,
    thetal_f90ptr phinh_i, thetal_f90ptr v, thetal_f90ptr w_i, thetal_f90ptr eta_dot_dpdn, thetal_f90ptr vn0)
{
  (void)eta_dot_dpdn; (void)vn0;
  init_elements_states_c(v, w_i, vtheta_dp, phinh_i, dp3d, nullptr, ps_v);
}

void init_reference_states_c(thetal_f90ptr theta_ref, thetal_f90ptr dp_ref, thetal_f90ptr phi_ref,
                             thetal_f90ptr nu_scale_top)
{
  REQUIRE(S.device_ready, "init_functors_c must precede init_reference_states_c");
  if (!S.device_ready) return;
  const size_t n = S.prm.nelem, nlev = S.nlev;
  CUDA_OK(cudaMemcpy((void *)S.ref.theta_ref, *theta_ref, n * 16 * nlev * sizeof(double), cudaMemcpyHostToDevice));
  CUDA_OK(cudaMemcpy((void *)S.ref.dp_ref, *dp_ref, n * 16 * nlev * sizeof(double), cudaMemcpyHostToDevice));
  CUDA_OK(cudaMemcpy((void *)S.ref.phi_ref, *phi_ref, n * 16 * (nlev + 1) * sizeof(double), cudaMemcpyHostToDevice));
  S.nu_scale_top.assign(*nu_scale_top, *nu_scale_top + nlev);
  int ntom = 0;                                     // model_init_mod.F90:128
  for (size_t k = 0; k < nlev; k++) if (S.nu_scale_top[k] > 0) ntom = (int)k + 1;
  S.prm.nlev_tom = ntom;
  upload_constants(nullptr, nullptr, nullptr, nullptr, S.nu_scale_top.data(), (int)nlev);
}

// theta_f2c_mod.F90:189-198, cxx_f90_interface_theta.cpp:231-262: derived%FM(np,np,3,nlev,nelemd),
// FVTheta(np,np,nlev,nelemd), FT (unused by the NH dynamics forcing), FPHI(np,np,nlevp,nelemd); FQ needs
// tracers (qsize = 0 here).  Applied by prim_run_subcycle_c when ftype is 0 or 2.
void push_forcing_to_c(const double *fm, const double *fvtheta, const double *ft, const double *fphi, const double *fq)
{
  (void)ft; (void)fq;
  REQUIRE(S.device_ready, "init_functors_c must precede push_forcing_to_c");
  REQUIRE(fm && fvtheta && fphi, "push_forcing_to_c: FM, FVTheta and FPHI are required");
  const size_t n = S.prm.nelem, nlev = S.nlev;
  if (!S.d_fm) {
    S.d_fm = dalloc<double>(n * 48 * nlev);
    S.d_fvtheta = dalloc<double>(n * 16 * nlev);
    S.d_fphi = dalloc<double>(n * 16 * (nlev + 1));
  }
  CUDA_OK(cudaMemcpyAsync(S.d_fm, fm, n * 48 * nlev * sizeof(double), cudaMemcpyHostToDevice, S.stream));
  CUDA_OK(cudaMemcpyAsync(S.d_fvtheta, fvtheta, n * 16 * nlev * sizeof(double), cudaMemcpyHostToDevice, S.stream));
  CUDA_OK(cudaMemcpyAsync(S.d_fphi, fphi, n * 16 * (nlev + 1) * sizeof(double), cudaMemcpyHostToDevice, S.stream));
  CUDA_OK(cudaStreamSynchronize(S.stream));
  S.have_forcing = true;
}

// applyCAMforcing_dynamics (PA:517-541, standalone HOMME: FM(3) forces w_i) on time level *nt (0-based)
void apply_cam_forcing_dynamics_c(const int *nt, const double *dt)
{
  REQUIRE(S.device_ready && *nt >= 0 && *nt < NTL, "apply_cam_forcing_dynamics_c: bad state/time level");
  REQUIRE(S.have_forcing, "apply_cam_forcing_dynamics_c: push_forcing_to_c has not been called");
  launch_forcing(S.tl[*nt], S.d_fm, S.d_fvtheta, S.d_fphi, S.geo, *dt, S.prm.nelem, S.nlev, S.stream);
  S.launches += 1;
  thetal_b200_sync();
}

// Addition: DSS of a caller's field, field(np,np,nlayers,nelemd) in place: every node ends with the sum
// over the elements that share it, in the Fortran unpack order (edge_mod.F90:888-941), across ranks included.
void dss_c(thetal_f90ptr_out field, const int *nlayers)
{
  REQUIRE(S.device_ready && field && *field, "dss_c before init_functors_c");
  const int nl = *nlayers, n = S.prm.nelem;
  REQUIRE(nl >= 1 && nl <= 6 * S.nlev + 1, "dss_c: nlayers must be in [1, 6*nlev+1] (size of the halo buffers)");
  const size_t own = (size_t)n * nl * 16, all = (size_t)(n + S.nghost) * nl * 16;
  double *d_f = nullptr, *d_o = nullptr;
  CUDA_OK(cudaMalloc(&d_f, all * sizeof(double)));
  CUDA_OK(cudaMalloc(&d_o, own * sizeof(double)));
  CUDA_OK(cudaMemcpyAsync(d_f, *field, own * sizeof(double), cudaMemcpyHostToDevice, S.stream));
  if (!S.peers.empty()) {
    launch_halo_generic(d_f, S.d_send_idx, (int)S.hsend_idx.size(), nl, S.d_sendbuf, 0, S.stream);
    NCCL_OK(ncclGroupStart());
    for (const Peer &p : S.peers) {
      NCCL_OK(ncclSend(S.d_sendbuf + (size_t)p.send_off * nl, (size_t)p.send_cnt * nl, ncclDouble, p.rank, S.comm, S.stream));
      NCCL_OK(ncclRecv(S.d_recvbuf + (size_t)p.recv_off * nl, (size_t)p.recv_cnt * nl, ncclDouble, p.rank, S.comm, S.stream));
    }
    NCCL_OK(ncclGroupEnd());
    launch_halo_generic(d_f, S.d_recv_idx, (int)S.hrecv_idx.size(), nl, S.d_recvbuf, 1, S.stream);
    S.launches += 2;
  }
  launch_dss_generic(d_f, d_o, S.geo.dss_src, n, nl, S.stream);
  S.launches += 1;
  CUDA_OK(cudaMemcpyAsync(*field, d_o, own * sizeof(double), cudaMemcpyDeviceToHost, S.stream));
  CUDA_OK(cudaStreamSynchronize(S.stream));
  cudaFree(d_f); cudaFree(d_o);
}

void sync_diagnostics_to_host_c(void)
{
  if (!S.device_ready) return;
  unsigned long long bits = 0; int it = 0;
  CUDA_OK(cudaMemcpyAsync(&bits, S.d_maxdelta, sizeof bits, cudaMemcpyDeviceToHost, S.stream));
  CUDA_OK(cudaMemcpyAsync(&it, S.d_maxiter, sizeof it, cudaMemcpyDeviceToHost, S.stream));
  CUDA_OK(cudaStreamSynchronize(S.stream));
  S.max_itercnt = it;
  memcpy(&S.max_deltaerr, &bits, sizeof bits);
}

void thetal_b200_get_derived(double *vn0, double *eta_dot_dpdn, double *omega_p)
{
  const size_t n = S.prm.nelem, nlev = S.nlev;
  if (vn0) CUDA_OK(cudaMemcpy(vn0, S.der.vn0, n * 32 * nlev * sizeof(double), cudaMemcpyDeviceToHost));
  if (eta_dot_dpdn) CUDA_OK(cudaMemcpy(eta_dot_dpdn, S.der.eta_dot_dpdn, n * 16 * (nlev + 1) * sizeof(double), cudaMemcpyDeviceToHost));
  if (omega_p) CUDA_OK(cudaMemcpy(omega_p, S.der.omega_p, n * 16 * nlev * sizeof(double), cudaMemcpyDeviceToHost));
}
void thetal_b200_zero_derived(void)
{
  const size_t n = S.prm.nelem, nlev = S.nlev;
  CUDA_OK(cudaMemsetAsync(S.der.vn0, 0, n * 32 * nlev * sizeof(double), S.stream));
  CUDA_OK(cudaMemsetAsync(S.der.eta_dot_dpdn, 0, n * 16 * (nlev + 1) * sizeof(double), S.stream));
  CUDA_OK(cudaMemsetAsync(S.der.omega_p, 0, n * 16 * nlev * sizeof(double), S.stream));
}
void thetal_b200_sync(void)
{
  if (!S.device_ready) return;
  CUDA_OK(cudaStreamSynchronize(S.stream));
  CUDA_OK(cudaGetLastError());
  check_device_errors("thetal_b200_sync");
}
long thetal_b200_launch_count(void) { return S.launches; }

// Measured FP64 FMA rate of this GPU in TFLOP/s (2 flops per DFMA): independent DFMA chains on every SM,
// timed with CUDA events on the library stream.  The second roofline axis of the FP64-bound kernels.
double thetal_b200_measure_fp64_peak(void)
{
  REQUIRE_R(S.active, "thetal_b200_measure_fp64_peak before initialize_hommexx_session", -1.0);
  double best = 0;
  for (int rep = 0; rep < 4; rep++) {
    CUDA_OK_R(cudaEventRecord(S.ev0, S.stream), -1.0);
    const double flops = launch_fp64_peak(S.stream);
    CUDA_OK_R(cudaEventRecord(S.ev1, S.stream), -1.0);
    CUDA_OK_R(cudaEventSynchronize(S.ev1), -1.0);
    float ms = 0;
    CUDA_OK_R(cudaEventElapsedTime(&ms, S.ev0, S.ev1), -1.0);
    if (rep > 0 && ms > 0) best = std::max(best, flops / (ms * 1e-3) / 1e12);
  }
  return best;
}

// Partition-boundary tables built by finalize_connectivity (host only, no device needed): used by the
// CPU tests of the multi-rank host logic and by drivers that want to report halo sizes.
void thetal_b200_halo_counts(int *nghost, int *npeers, int *nsend, int *nrecv)
{
  *nghost = S.nghost; *npeers = (int)S.peers.size();
  *nsend = (int)S.hsend_idx.size(); *nrecv = (int)S.hrecv_idx.size();
}
void thetal_b200_halo_lists(int *peer_rank, int *send_off, int *send_cnt, int *recv_off, int *recv_cnt,
                            int *send_idx, int *recv_idx)
{
  for (size_t i = 0; i < S.peers.size(); i++) {
    peer_rank[i] = S.peers[i].rank; send_off[i] = S.peers[i].send_off; send_cnt[i] = S.peers[i].send_cnt;
    recv_off[i] = S.peers[i].recv_off; recv_cnt[i] = S.peers[i].recv_cnt;
  }
  memcpy(send_idx, S.hsend_idx.data(), S.hsend_idx.size() * sizeof(int));
  memcpy(recv_idx, S.hrecv_idx.data(), S.hrecv_idx.size() * sizeof(int));
}
void thetal_b200_dss_table(int *dss_src) { memcpy(dss_src, S.hdss.data(), S.hdss.size() * sizeof(int)); }
void thetal_b200_halo_stats(long *exchanges, long *bytes) { *exchanges = S.exchanges; *bytes = S.halo_bytes; }

// Global diagnostics of time level *tl on the device: prim_energy_halftimes
// (prim_state_mod.F90:822-939) + global_integral (global_norms_mod.F90:47-90, divided by the 4 pi
// domain size) with an order-independent sum, so every rank of every partition gets the same bits.
//   out[0..7] = min ps, max ps, <ps>, <KE>, <PE>, <IE>, <KE+PE+IE>, elements owned by this rank
void thetal_b200_global_diagnostics(const int *tl, double *out)
{
  REQUIRE(S.device_ready && *tl >= 0 && *tl < NTL, "thetal_b200_global_diagnostics: bad state/time level");
  constexpr int NL = thetal_reprosum::NLIMB, NACC = NDIAG + 2 + NDIAG * NL;
  unsigned long long *maxbits = S.d_diag_acc, *psmm = S.d_diag_acc + NDIAG, *limbs = S.d_diag_acc + NDIAG + 2;
  CUDA_OK(cudaMemsetAsync(S.d_diag_acc, 0, NACC * sizeof(unsigned long long), S.stream));
  CUDA_OK(cudaMemsetAsync(psmm, 0xff, sizeof(unsigned long long), S.stream));          // running min
  DiagArgs a;
  a.st = S.tl[*tl]; a.geo = S.geo; a.prm = S.prm; a.part = S.d_diag_part; a.maxbits = maxbits; a.ps_minmax = psmm;
  launch_diag_elements(a, S.stream); S.launches += 1;
  if (S.nranks > 1) {
    NCCL_OK(ncclGroupStart());
    NCCL_OK(ncclAllReduce(maxbits, maxbits, NDIAG, ncclUint64, ncclMax, S.comm, S.stream));
    NCCL_OK(ncclAllReduce(psmm, psmm, 1, ncclUint64, ncclMin, S.comm, S.stream));
    NCCL_OK(ncclAllReduce(psmm + 1, psmm + 1, 1, ncclUint64, ncclMax, S.comm, S.stream));
    NCCL_OK(ncclGroupEnd());
  }
  launch_diag_accumulate(S.d_diag_part, S.prm.nelem, maxbits, limbs, S.stream); S.launches += 1;
  if (S.nranks > 1) NCCL_OK(ncclAllReduce(limbs, limbs, NDIAG * NL, ncclInt64, ncclSum, S.comm, S.stream));
  unsigned long long h[NACC];
  CUDA_OK(cudaMemcpyAsync(h, S.d_diag_acc, sizeof h, cudaMemcpyDeviceToHost, S.stream));
  CUDA_OK(cudaStreamSynchronize(S.stream));
  check_device_errors("thetal_b200_global_diagnostics");
  memcpy(&out[0], &h[NDIAG], 8); memcpy(&out[1], &h[NDIAG + 1], 8);
  const double four_pi = 4.0 * 3.14159265358979323846264338327950288;       // dd_pi, physical_constants.F90:48
  for (int q = 0; q < NDIAG; q++) {
    long long l[NL];
    for (int j = 0; j < NL; j++) l[j] = (long long)h[NDIAG + 2 + q * NL + j];
    out[2 + q] = thetal_reprosum::combine(l, thetal_reprosum::exponent_of_bits(h[q])) / four_pi;
  }
  out[6] = out[3] + out[4] + out[5];
  out[7] = (double)S.prm.nelem;      // elements owned by this rank
}

// thetal_b200_global_diagnostics for callers that may not allocate (graph capture, harnesses that forbid
// cudaMalloc): the same two kernels, with the element integrals and the accumulators in file-scope device
// arrays of a fixed maximum size instead of the session's buffers.  No counterpart in the reference.  Every
// GPU that loads the library holds the arrays (THETAL_B200_NOALLOC_MAX_ELEMS * 32 B + 144 B = 512 KB) until
// exit; all calls share them, so calls that overlap on the GPU (two host threads, two streams) race.  The
// accumulators are cleared on every call; the element integrals are fully rewritten for elements < nelem.
// In host code the array names below are stand-ins, not device addresses: the address a kernel can use comes
// from cudaGetSymbolAddress, resolved once per device with that device current.
constexpr int NOALLOC_NACC = NDIAG + 2 + NDIAG * thetal_reprosum::NLIMB;
static __device__ __align__(256) double g_noalloc_part[(size_t)THETAL_B200_NOALLOC_MAX_ELEMS * NDIAG];
static __device__ __align__(256) unsigned long long g_noalloc_acc[NOALLOC_NACC];
constexpr int NOALLOC_MAX_DEVICES = 64;
static double *g_noalloc_part_addr[NOALLOC_MAX_DEVICES];
static unsigned long long *g_noalloc_acc_addr[NOALLOC_MAX_DEVICES];

void thetal_b200_global_diagnostics_noalloc(const int *tl, double *out)
{
  REQUIRE(S.device_ready && *tl >= 0 && *tl < NTL, "thetal_b200_global_diagnostics_noalloc: bad state/time level");
  REQUIRE(S.prm.nelem <= THETAL_B200_NOALLOC_MAX_ELEMS,
          "thetal_b200_global_diagnostics_noalloc: " + std::to_string(S.prm.nelem) + " elements on this rank exceed the "
          "static scratch of THETAL_B200_NOALLOC_MAX_ELEMS = " + std::to_string(THETAL_B200_NOALLOC_MAX_ELEMS) +
          "; use thetal_b200_global_diagnostics");
  REQUIRE(S.nranks == 1, "thetal_b200_global_diagnostics_noalloc: single-rank sessions only (a first NCCL collective "
                         "may allocate); use thetal_b200_global_diagnostics");
  int dev = -1;
  CUDA_OK(cudaGetDevice(&dev));
  REQUIRE(dev >= 0 && dev < NOALLOC_MAX_DEVICES, "thetal_b200_global_diagnostics_noalloc: device ordinal out of range");
  if (!g_noalloc_part_addr[dev]) {       // set-up, once per device, the session's device current
    double *pp = nullptr; unsigned long long *pa = nullptr;
    CUDA_OK(cudaGetSymbolAddress((void **)&pp, g_noalloc_part));
    CUDA_OK(cudaGetSymbolAddress((void **)&pa, g_noalloc_acc));
    REQUIRE(pp && pa, "thetal_b200_global_diagnostics_noalloc: cudaGetSymbolAddress returned a null address");
    g_noalloc_part_addr[dev] = pp; g_noalloc_acc_addr[dev] = pa;
  }
  double *part = g_noalloc_part_addr[dev];
  unsigned long long *acc = g_noalloc_acc_addr[dev];
  constexpr int NL = thetal_reprosum::NLIMB, NACC = NOALLOC_NACC;
  unsigned long long *maxbits = acc, *psmm = acc + NDIAG, *limbs = acc + NDIAG + 2;
  CUDA_OK(cudaMemsetAsync(acc, 0, NACC * sizeof(unsigned long long), S.stream));      // the last call's contents
  CUDA_OK(cudaMemsetAsync(psmm, 0xff, sizeof(unsigned long long), S.stream));          // running min
  DiagArgs a;
  a.st = S.tl[*tl]; a.geo = S.geo; a.prm = S.prm; a.part = part; a.maxbits = maxbits; a.ps_minmax = psmm;
  launch_diag_elements(a, S.stream); S.launches += 1;
  launch_diag_accumulate(part, S.prm.nelem, maxbits, limbs, S.stream); S.launches += 1;
  unsigned long long h[NACC];
  CUDA_OK(cudaMemcpyAsync(h, acc, sizeof h, cudaMemcpyDeviceToHost, S.stream));
  CUDA_OK(cudaStreamSynchronize(S.stream));
  check_device_errors("thetal_b200_global_diagnostics_noalloc");
  memcpy(&out[0], &h[NDIAG], 8); memcpy(&out[1], &h[NDIAG + 1], 8);
  const double four_pi = 4.0 * 3.14159265358979323846264338327950288;       // dd_pi, physical_constants.F90:48
  for (int q = 0; q < NDIAG; q++) {
    long long l[NL];
    for (int j = 0; j < NL; j++) l[j] = (long long)h[NDIAG + 2 + q * NL + j];
    out[2 + q] = thetal_reprosum::combine(l, thetal_reprosum::exponent_of_bits(h[q])) / four_pi;
  }
  out[6] = out[3] + out[4] + out[5];
  out[7] = (double)S.prm.nelem;
}

void thetal_b200_profile_step(const double *dt, const int *nm1, const int *n0, const int *np1,
                              double *ms_by_family, int nfamily);

// Field extrema of time level *tl as prim_printstate reports them (prim_state_mod.F90:219-330): out[2f], out[2f+1] =
// global min, max of f = u, v, w_i, theta, T, dz, dp3d (7 fields, 14 numbers), the same on every rank.
void thetal_b200_field_extrema(const int *tl, double *out)
{
  REQUIRE(S.device_ready && *tl >= 0 && *tl < NTL, "thetal_b200_field_extrema: bad state/time level");
  if (!S.d_extrema) S.d_extrema = dalloc<unsigned long long>(2 * NEXTREMA);     // freed with the session
  unsigned long long *d_keys = S.d_extrema;
  unsigned long long init[2 * NEXTREMA], h[2 * NEXTREMA];
  for (int f = 0; f < NEXTREMA; f++) { init[2 * f] = ~0ULL; init[2 * f + 1] = 0ULL; }
  CUDA_OK(cudaMemcpyAsync(d_keys, init, sizeof init, cudaMemcpyHostToDevice, S.stream));
  DiagArgs a;
  a.st = S.tl[*tl]; a.geo = S.geo; a.prm = S.prm; a.part = nullptr; a.maxbits = nullptr; a.ps_minmax = nullptr;
  launch_extrema(a, d_keys, S.stream); S.launches += 1;
  CUDA_OK(cudaMemcpyAsync(h, d_keys, sizeof h, cudaMemcpyDeviceToHost, S.stream));
  CUDA_OK(cudaStreamSynchronize(S.stream));
  if (S.nranks > 1) {       // min and max keys of all ranks: interleaved, so two strided reductions on the host copy
    unsigned long long mins[NEXTREMA], maxs[NEXTREMA];
    for (int f = 0; f < NEXTREMA; f++) { mins[f] = h[2 * f]; maxs[f] = h[2 * f + 1]; }
    CUDA_OK(cudaMemcpyAsync(d_keys, mins, sizeof mins, cudaMemcpyHostToDevice, S.stream));
    CUDA_OK(cudaMemcpyAsync(d_keys + NEXTREMA, maxs, sizeof maxs, cudaMemcpyHostToDevice, S.stream));
    NCCL_OK(ncclGroupStart());
    NCCL_OK(ncclAllReduce(d_keys, d_keys, NEXTREMA, ncclUint64, ncclMin, S.comm, S.stream));
    NCCL_OK(ncclAllReduce(d_keys + NEXTREMA, d_keys + NEXTREMA, NEXTREMA, ncclUint64, ncclMax, S.comm, S.stream));
    NCCL_OK(ncclGroupEnd());
    CUDA_OK(cudaMemcpyAsync(mins, d_keys, sizeof mins, cudaMemcpyDeviceToHost, S.stream));
    CUDA_OK(cudaMemcpyAsync(maxs, d_keys + NEXTREMA, sizeof maxs, cudaMemcpyDeviceToHost, S.stream));
    CUDA_OK(cudaStreamSynchronize(S.stream));
    for (int f = 0; f < NEXTREMA; f++) { h[2 * f] = mins[f]; h[2 * f + 1] = maxs[f]; }
  }
  for (int i = 0; i < 2 * NEXTREMA; i++) {
    const unsigned long long k = h[i], b = (k >> 63) ? (k & 0x7fffffffffffffffULL) : ~k;
    memcpy(&out[i], &b, 8);
  }
}


// ---------------------------------------------------------------- the step
static bool tl_ok(int t) { return t >= 0 && t < NTL; }

void caar_c(const int *np1, const int *nm1, const int *n0, const double *dt2, const double *eta_ave_w,
            const double *scale1, const double *scale2, const double *scale3)
{
  REQUIRE(S.device_ready && tl_ok(*np1) && tl_ok(*nm1) && tl_ok(*n0), "caar_c: bad state/time level");
  run_caar(*np1, *nm1, *n0, *dt2, *eta_ave_w, *scale1, *scale2, *scale3);
  thetal_b200_sync();
}

void dirk_c(const int *nm1, const double *alphadt_nm1, const int *n0, const double *alphadt_n0,
            const int *np1, const double *dt2, int *itercount, double *itererr)
{
  REQUIRE(S.device_ready && tl_ok(*np1) && tl_ok(*nm1) && tl_ok(*n0), "dirk_c: bad state/time level");
  CUDA_OK(cudaMemsetAsync(S.d_maxiter, 0, sizeof(int), S.stream));
  CUDA_OK(cudaMemsetAsync(S.d_maxdelta, 0, sizeof(unsigned long long), S.stream));
  run_dirk(*nm1, *alphadt_nm1, *n0, *alphadt_n0, *np1, *dt2);
  thetal_b200_sync();
  sync_diagnostics_to_host_c();
  if (itercount) *itercount = S.max_itercnt;
  if (itererr) *itererr = S.max_deltaerr;
}

// Device time of compute_stage_value_dirk alone (tools/bench_dirk.py): w_i and phinh_i of np1 are saved and
// restored before every repetition, so each launch solves the same stage problem; returns ms per launch.
double thetal_b200_time_dirk(const int *nm1, const double *alphadt_nm1, const int *n0, const double *alphadt_n0,
                             const int *np1, const double *dt2, const int *reps, int *itercount)
{
  REQUIRE_R(S.device_ready && tl_ok(*np1) && tl_ok(*nm1) && tl_ok(*n0), "thetal_b200_time_dirk: bad state/time level", -1.0);
  const size_t bytes = (size_t)S.prm.nelem * 16 * (S.nlev + 1) * sizeof(double);
  const StatePtrs &st = S.tl[*np1];
  CUDA_OK_R(cudaMemcpyAsync(S.tensB.w, st.w, bytes, cudaMemcpyDeviceToDevice, S.stream), -1.0);
  CUDA_OK_R(cudaMemcpyAsync(S.tensB.phi, st.phi, bytes, cudaMemcpyDeviceToDevice, S.stream), -1.0);
  CUDA_OK_R(cudaMemsetAsync(S.d_maxiter, 0, sizeof(int), S.stream), -1.0);
  double total = 0;
  for (int r = 0; r < *reps; r++) {
    CUDA_OK_R(cudaMemcpyAsync(st.w, S.tensB.w, bytes, cudaMemcpyDeviceToDevice, S.stream), -1.0);
    CUDA_OK_R(cudaMemcpyAsync(st.phi, S.tensB.phi, bytes, cudaMemcpyDeviceToDevice, S.stream), -1.0);
    CUDA_OK_R(cudaEventRecord(S.ev0, S.stream), -1.0);
    run_dirk(*nm1, *alphadt_nm1, *n0, *alphadt_n0, *np1, *dt2);
    CUDA_OK_R(cudaEventRecord(S.ev1, S.stream), -1.0);
    CUDA_OK_R(cudaEventSynchronize(S.ev1), -1.0);
    float ms = 0;
    CUDA_OK_R(cudaEventElapsedTime(&ms, S.ev0, S.ev1), -1.0);
    total += ms;
  }
  sync_diagnostics_to_host_c();
  if (itercount) *itercount = S.max_itercnt;
  check_device_errors("thetal_b200_time_dirk");
  return total / (*reps > 0 ? *reps : 1);
}

void advance_hypervis_c(const int *nt, const double *dt2, const double *eta_ave_w)
{
  REQUIRE(S.device_ready && tl_ok(*nt), "advance_hypervis_c: bad state/time level");
  run_hypervis(*nt, *dt2, *eta_ave_w);
  thetal_b200_sync();
}

void prim_advance_exp_c(const double *dt, const int *nm1, const int *n0, const int *np1,
                        const int *nstep, const int *compute_diagnostics)
{
  (void)nstep; (void)compute_diagnostics;
  REQUIRE(S.device_ready && tl_ok(*np1) && tl_ok(*nm1) && tl_ok(*n0), "prim_advance_exp_c: bad state/time level");
  run_prim_advance_exp(*dt, *nm1, *n0, *np1);
  thetal_b200_sync();
}

// vertical_remap (vertremap_mod.F90:23-171) of time level *np1 (0-based): back from the
// floating Lagrangian levels to the reference levels with PPM (vert_remap_u_alg = remap_alg 1/10/11).
static void run_vertical_remap(int np1)
{
  RemapArgs a;
  a.st = S.tl[np1]; a.geo = S.geo; a.prm = S.prm; a.remap_alg = S.remap_alg; a.err = S.d_err;
  launch_vertical_remap(a, S.stream); S.launches += 1;
}
void vertical_remap_c(const int *np1)
{
  REQUIRE(S.device_ready && tl_ok(*np1), "vertical_remap_c: bad state/time level");
  run_vertical_remap(*np1);
  thetal_b200_sync();
}
void thetal_b200_set_vertical_remap(const int *on) { S.do_remap = *on; }

// prim_run_subcycle (prim_driver_base.F90:1024-1195, prim_driver.cpp:60-181) restricted to the
// dynamics (qsize = 0): rsplit x prim_step, each prim_step = zero the derived accumulators
// (set_tracer_transport_derived_values, prim_step.cpp:24-66) + dt_tracer_factor dynamics steps with
// a leapfrog time-level update between them; then vertical_remap on np1 (prim_driver.cpp:143-151) and
// the last TimeLevel update (:170).  nstep, nm1, n0, np1 are pure OUTPUTS, 0-based, taken from the
// library's own TimeLevel (set by init_time_level_c) exactly as prim_driver.cpp:174-178 does; the
// Fortran caller passes uninitialised integers and adds 1 afterwards (prim_driver_mod.F90:554-560).
static void leapfrog_update()
{ // TimeLevel::update_dynamics_levels(LEAPFROG), time_mod.F90:109-138
  const int tmp = S.np1; S.np1 = S.nm1; S.nm1 = S.n0; S.n0 = tmp; S.nstep++;
}
static void zero_derived_async()
{
  const size_t n = S.prm.nelem, nlev = S.nlev;
  cudaMemsetAsync(S.der.vn0, 0, n * 32 * nlev * sizeof(double), S.stream);
  cudaMemsetAsync(S.der.eta_dot_dpdn, 0, n * 16 * (nlev + 1) * sizeof(double), S.stream);
  cudaMemsetAsync(S.der.omega_p, 0, n * 16 * nlev * sizeof(double), S.stream);
}
static void run_prim_step(double dt)
{ // prim_step.cpp:68-117
  zero_derived_async();
  run_prim_advance_exp(dt, S.nm1, S.n0, S.np1);
  for (int n = 1; n < S.prm.qsplit; n++) {
    leapfrog_update();
    run_prim_advance_exp(dt, S.nm1, S.n0, S.np1);
  }
}
void prim_run_subcycle_c(const double *tstep, int *nstep, int *nm1, int *n0, int *np1,
                         const int *next_output_step, const int *nsplit_iter)
{
  (void)next_output_step; (void)nsplit_iter;
  REQUIRE(S.device_ready, "prim_run_subcycle_c before init");
  REQUIRE(tl_ok(S.nm1) && tl_ok(S.n0) && tl_ok(S.np1) && S.nm1 != S.n0 && S.n0 != S.np1 && S.nm1 != S.np1,
          "prim_run_subcycle_c: init_time_level_c did not set three distinct time levels");
  // prim_driver.cpp:75-83: dt_remap = dt*dt_remap_factor, or dt*dt_tracer_factor in the Eulerian case
  const double dt_remap = *tstep * (S.prm.rsplit > 0 ? S.prm.rsplit * S.prm.qsplit : S.prm.qsplit);
  if ((S.ftype == 0 || S.ftype == 2) && S.have_forcing) {                  // prim_driver.cpp:117-125
    launch_forcing(S.tl[S.n0], S.d_fm, S.d_fvtheta, S.d_fphi, S.geo, dt_remap, S.prm.nelem, S.nlev, S.stream);
    S.launches += 1;
  }
  run_prim_step(*tstep);
  for (int r = 1; r < S.prm.rsplit; r++) {
    leapfrog_update();
    run_prim_step(*tstep);
  }
  if (S.do_remap && S.prm.rsplit > 0) run_vertical_remap(S.np1);          // rsplit = 0: tracers only (none)
  thetal_b200_sync();
  leapfrog_update();
  *nstep = S.nstep; *nm1 = S.nm1; *n0 = S.n0; *np1 = S.np1;
}

double thetal_b200_time_steps(const double *dt, const int *nsteps, int *nm1, int *n0, int *np1)
{
  REQUIRE_R(S.device_ready, "thetal_b200_time_steps before init", -1.0);
  int t_nm1 = *nm1, t_n0 = *n0, t_np1 = *np1;
  CUDA_OK_R(cudaEventRecord(S.ev0, S.stream), -1.0);
  for (int s = 0; s < *nsteps; s++) {
    run_prim_advance_exp(*dt, t_nm1, t_n0, t_np1);
    const int tmp = t_np1; t_np1 = t_nm1; t_nm1 = t_n0; t_n0 = tmp;
  }
  CUDA_OK_R(cudaEventRecord(S.ev1, S.stream), -1.0);
  CUDA_OK_R(cudaEventSynchronize(S.ev1), -1.0);
  float ms = 0;
  CUDA_OK_R(cudaEventElapsedTime(&ms, S.ev0, S.ev1), -1.0);
  *nm1 = t_nm1; *n0 = t_n0; *np1 = t_np1;
  check_device_errors("thetal_b200_time_steps");
  return (double)ms;
}

}  // extern "C"

namespace { void run_prim_advance_exp(double dt, int nm1, int n0, int np1); }
extern "C" void thetal_b200_profile_step(const double *dt, const int *nm1, const int *n0, const int *np1,
                                         double *ms_by_family, int nfamily)
{
  REQUIRE(S.device_ready, "thetal_b200_profile_step before init");
  if (!S.pe0) { CUDA_OK(cudaEventCreate(&S.pe0)); CUDA_OK(cudaEventCreate(&S.pe1)); }
  for (double &m : S.prof_ms) m = 0;
  CUDA_OK(cudaStreamSynchronize(S.stream));
  S.prof = true;
  run_prim_advance_exp(*dt, *nm1, *n0, *np1);
  S.prof = false;
  CUDA_OK(cudaStreamSynchronize(S.stream));
  for (int i = 0; i < nfamily && i < 8; i++) ms_by_family[i] = S.prof_ms[i];
}

// ---- single-time-level transfers (bench e2e; a driver that only needs the new state) ----------
extern "C" void thetal_b200_push_level(const int *tl, const double *v, const double *w_i, const double *vtheta_dp,
                                       const double *phinh_i, const double *dp3d)
{ // host arrays [ie][k][(c)][16] of ONE time level -> device, asynchronous on the library stream
  REQUIRE(S.device_ready && *tl >= 0 && *tl < NTL, "thetal_b200_push_level");
  const size_t n = S.prm.nelem, nlev = S.nlev;
  const StatePtrs &d = S.tl[*tl];
  CUDA_OK(cudaMemcpyAsync(d.v, v, n * 32 * nlev * sizeof(double), cudaMemcpyHostToDevice, S.stream));
  CUDA_OK(cudaMemcpyAsync(d.w, w_i, n * 16 * (nlev + 1) * sizeof(double), cudaMemcpyHostToDevice, S.stream));
  CUDA_OK(cudaMemcpyAsync(d.th, vtheta_dp, n * 16 * nlev * sizeof(double), cudaMemcpyHostToDevice, S.stream));
  CUDA_OK(cudaMemcpyAsync(d.phi, phinh_i, n * 16 * (nlev + 1) * sizeof(double), cudaMemcpyHostToDevice, S.stream));
  CUDA_OK(cudaMemcpyAsync(d.dp, dp3d, n * 16 * nlev * sizeof(double), cudaMemcpyHostToDevice, S.stream));
}
extern "C" void thetal_b200_pull_level(const int *tl, double *v, double *w_i, double *vtheta_dp,
                                       double *phinh_i, double *dp3d)
{
  REQUIRE(S.device_ready && *tl >= 0 && *tl < NTL, "thetal_b200_pull_level");
  const size_t n = S.prm.nelem, nlev = S.nlev;
  const StatePtrs &d = S.tl[*tl];
  CUDA_OK(cudaMemcpyAsync(v, d.v, n * 32 * nlev * sizeof(double), cudaMemcpyDeviceToHost, S.stream));
  CUDA_OK(cudaMemcpyAsync(w_i, d.w, n * 16 * (nlev + 1) * sizeof(double), cudaMemcpyDeviceToHost, S.stream));
  CUDA_OK(cudaMemcpyAsync(vtheta_dp, d.th, n * 16 * nlev * sizeof(double), cudaMemcpyDeviceToHost, S.stream));
  CUDA_OK(cudaMemcpyAsync(phinh_i, d.phi, n * 16 * (nlev + 1) * sizeof(double), cudaMemcpyDeviceToHost, S.stream));
  CUDA_OK(cudaMemcpyAsync(dp3d, d.dp, n * 16 * nlev * sizeof(double), cudaMemcpyDeviceToHost, S.stream));
  CUDA_OK(cudaStreamSynchronize(S.stream));
}
extern "C" void thetal_b200_pin_host(void *ptr, const long *bytes, const int *pin)
{
  if (*pin) CUDA_OK(cudaHostRegister(ptr, (size_t)*bytes, cudaHostRegisterDefault));
  else CUDA_OK(cudaHostUnregister(ptr));
}
// ---- pipelined single-step service --------------------------------------------------------------
// One dynamics step per call on an INDEPENDENT input: the n0 state comes from pinned host arrays
// ([ie][k][(c)][16] of one time level), the np1 state goes to pinned host arrays.  The host->device
// copy of call j+1, the kernels of call j and the device->host copy of call j-1 run on three streams;
// the staging buffers change places with the time-level slots by pointer (no device copies).  The
// device state is scratch while the pipeline is in use (slots 0,1,2 = nm1,n0,np1); outputs are
// complete after thetal_b200_pipeline_flush().
namespace {
void pipeline_setup()
{
  const int n = S.prm.nelem, nlev = S.nlev;
  for (int b = 0; b < 2; b++) {
    S.pipe_in[b] = alloc_state(n, nlev); S.pipe_out[b] = alloc_state(n, nlev);
    // the step never writes phinh_i at the surface: every buffer that can become slot np1 carries phis there,
    // as every time level of the reference does (element_state.F90: phinh_i(:,:,nlevp,:) = phis)
    for (double *phi : {S.pipe_in[b].phi, S.pipe_out[b].phi})
      CUDA_OK(cudaMemcpy2DAsync(phi + (size_t)nlev * 16, 16 * (size_t)(nlev + 1) * sizeof(double), S.geo.phis,
                                16 * sizeof(double), 16 * sizeof(double), n, cudaMemcpyDeviceToDevice, S.stream));
  }
  CUDA_OK(cudaStreamCreateWithFlags(&S.s_h2d, cudaStreamNonBlocking));
  CUDA_OK(cudaStreamCreateWithFlags(&S.s_d2h, cudaStreamNonBlocking));
  for (int i = 0; i < 4; i++) {
    CUDA_OK(cudaEventCreateWithFlags(&S.ev_pin[i], cudaEventDisableTiming));
    CUDA_OK(cudaEventCreateWithFlags(&S.ev_pdone[i], cudaEventDisableTiming));
    CUDA_OK(cudaEventCreateWithFlags(&S.ev_pd2h[i], cudaEventDisableTiming));
  }
  S.pipe_calls = 0; S.pipe_ready = true;
}
void copy_level(const StatePtrs &d, const double *v, const double *w, const double *th, const double *phi,
                const double *dp, bool to_device, cudaStream_t st)
{
  const size_t n = S.prm.nelem, nlev = S.nlev;
  const cudaMemcpyKind kind = to_device ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost;
  const size_t bytes[5] = {n * 32 * nlev * 8, n * 16 * (nlev + 1) * 8, n * 16 * nlev * 8, n * 16 * (nlev + 1) * 8, n * 16 * nlev * 8};
  double *dev[5] = {d.v, d.w, d.th, d.phi, d.dp};
  const double *host[5] = {v, w, th, phi, dp};
  for (int i = 0; i < 5; i++) {
    if (to_device) CUDA_OK(cudaMemcpyAsync(dev[i], host[i], bytes[i], kind, st));
    else CUDA_OK(cudaMemcpyAsync(const_cast<double *>(host[i]), dev[i], bytes[i], kind, st));
  }
}
}  // namespace

extern "C" void thetal_b200_pipeline_step(const double *dt, const double *v_in, const double *w_in,
    const double *th_in, const double *phi_in, const double *dp_in,
    double *v_out, double *w_out, double *th_out, double *phi_out, double *dp_out)
{
  REQUIRE(S.device_ready, "thetal_b200_pipeline_step before init");
  if (!S.pipe_ready) pipeline_setup();
  const long j = S.pipe_calls++;
  const int b = (int)(j & 1), e = (int)(j & 3), e3 = (int)((j + 1) & 3);     // e3 = (j-3) mod 4
  // 1. input: the staging buffer was the n0 slot of call j-3
  if (j >= 3) CUDA_OK(cudaStreamWaitEvent(S.s_h2d, S.ev_pdone[e3], 0));
  copy_level(S.pipe_in[b], v_in, w_in, th_in, phi_in, dp_in, true, S.s_h2d);
  CUDA_OK(cudaEventRecord(S.ev_pin[e], S.s_h2d));
  // 2. the step: staged input becomes slot n0; slot np1 is the buffer call j-3 copied out of
  CUDA_OK(cudaStreamWaitEvent(S.stream, S.ev_pin[e], 0));
  if (j >= 3) CUDA_OK(cudaStreamWaitEvent(S.stream, S.ev_pd2h[e3], 0));
  std::swap(S.tl[1], S.pipe_in[b]);
  run_prim_advance_exp(*dt, 0, 1, 2);
  std::swap(S.tl[2], S.pipe_out[b]);
  CUDA_OK(cudaEventRecord(S.ev_pdone[e], S.stream));
  // 3. output
  CUDA_OK(cudaStreamWaitEvent(S.s_d2h, S.ev_pdone[e], 0));
  copy_level(S.pipe_out[b], v_out, w_out, th_out, phi_out, dp_out, false, S.s_d2h);
  CUDA_OK(cudaEventRecord(S.ev_pd2h[e], S.s_d2h));
}
extern "C" void thetal_b200_pipeline_flush(void)
{
  if (!S.pipe_ready) return;
  CUDA_OK(cudaStreamSynchronize(S.s_h2d));
  CUDA_OK(cudaStreamSynchronize(S.stream));
  CUDA_OK(cudaStreamSynchronize(S.s_d2h));
  check_device_errors("thetal_b200_pipeline_flush");
}

// enqueue one dynamics step without synchronising (bench e2e)
extern "C" void thetal_b200_enqueue_step(const double *dt, const int *nm1, const int *n0, const int *np1)
{
  REQUIRE(S.device_ready, "thetal_b200_enqueue_step before init");
  run_prim_advance_exp(*dt, *nm1, *n0, *np1);
}
