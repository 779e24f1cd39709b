// include/shim/FluidSolverGPU.hpp -- header-only C++ host layer that re-creates the call surface of
// gidonbauer/FluidSolver's per-timestep path on top of the C ABI of include/fsgpu.h.
//
// A reference driver (scaling/LidDrivenFlow.cpp, test/TaylorGreenVortex.cpp, test/LaminarChannel.cpp)
// includes FS.hpp / Operators.hpp / BoundaryConditions.hpp / LinearSolver_StructHypre.hpp and calls
// free function templates on an `FS` object.  This header declares the same names with the same
// argument meaning; behind them every operation is a CUDA kernel of libfsgpu.so (no CPU fallback).
// INTEGRATION.md lists, call by call, what a driver changes.
//
// Differences a driver author must know (they follow from the state living in HBM):
//  * Field2D::operator()(i, j) reads / writes a HOST mirror.  After initialising fields with host
//    loops call fs.to_device(); before inspecting results call fs.to_host() (or field.download()).
//  * the loops the reference drivers inline (velocity update, pressure correction) are the calls
//    update_velocity(dU, dV, dt, fs) (exists in the reference: src/FS.hpp:423-436) and
//    correct_velocity(fs, delta_p, dt) (scaling/LidDrivenFlow.cpp:153-160).
//  * scratch fields (div, drhoUdt, drhoVdt, delta_p) are bound to the context's device fields the
//    first time they are passed to a call in that role.
#ifndef FSGPU_SHIM_FLUIDSOLVERGPU_HPP_
#define FSGPU_SHIM_FLUIDSOLVERGPU_HPP_

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <array>
#include <cmath>
#include <filesystem>
#include <source_location>
#include <string>
#include <string_view>
#include <utility>
#include <variant>
#include <vector>

#include "../fsgpu.h"

using Index = std::int64_t;  // src/Container.hpp

namespace fsgpu_shim {
// the reference's error convention: Igor::Panic -> message + exit(1) (Igor/Logging.hpp:50-62)
inline void check(int rc, const char* what) {
  if (rc != 0) {
    std::fprintf(stderr, "[fsgpu] %s failed: %s\n", what, fsgpu_last_error());
    std::exit(1);
  }
}
}  // namespace fsgpu_shim
#define FSGPU_CHECK(call) ::fsgpu_shim::check((call), #call)

// ---- Field2D / Field1D (src/Container.hpp:25-169) -------------------------------------------------
template <typename Float, Index N, Index NGHOST>
class Field1D {
  std::vector<Float> m_data = std::vector<Float>(static_cast<size_t>(N + 2 * NGHOST), Float{});
 public:
  Float& operator()(Index i) { return m_data[static_cast<size_t>(i + NGHOST)]; }
  const Float& operator()(Index i) const { return m_data[static_cast<size_t>(i + NGHOST)]; }
  Float* get_data() { return m_data.data(); }
  const Float* get_data() const { return m_data.data(); }
  static constexpr Index extent(Index) { return N; }
  static constexpr Index size() { return N + 2 * NGHOST; }
};

template <typename Float, Index M, Index N, Index NGHOST>
class Field2D {
  static_assert(sizeof(Float) == sizeof(double), "libfsgpu.so computes in FP64");
  std::vector<Float> m_data = std::vector<Float>(static_cast<size_t>((M + 2 * NGHOST) * (N + 2 * NGHOST)), Float{});
  fsgpu_ctx* m_ctx = nullptr;
  int m_field      = -1;
 public:
  Field2D() = default;
  Field2D(const Field2D&)            = delete;  // as the reference: non-copyable, non-movable
  Field2D& operator=(const Field2D&) = delete;
  // host mirror, Layout::C with j fastest (src/Container.hpp:105-111)
  Float& operator()(Index i, Index j) {
    return m_data[static_cast<size_t>((j + NGHOST) + (i + NGHOST) * (N + 2 * NGHOST))];
  }
  const Float& operator()(Index i, Index j) const {
    return m_data[static_cast<size_t>((j + NGHOST) + (i + NGHOST) * (N + 2 * NGHOST))];
  }
  Float* get_data() { return m_data.data(); }
  const Float* get_data() const { return m_data.data(); }
  static constexpr Index extent(Index d) { return d == 0 ? M : N; }
  static constexpr Index size() { return (M + 2 * NGHOST) * (N + 2 * NGHOST); }
  static constexpr Index n_ghost() { return NGHOST; }
  static constexpr bool is_valid_interior_index(Index i, Index j) { return i >= 0 && i < M && j >= 0 && j < N; }
  static constexpr bool is_valid_index(Index i, Index j) {
    return i >= -NGHOST && i < M + NGHOST && j >= -NGHOST && j < N + NGHOST;
  }
  // device binding
  void bind(fsgpu_ctx* ctx, int field) { m_ctx = ctx; m_field = field; }
  bool bound() const { return m_ctx != nullptr; }
  fsgpu_ctx* context() const { return m_ctx; }
  int field_id() const { return m_field; }
  void upload() const { if (m_ctx) FSGPU_CHECK(fsgpu_upload(m_ctx, m_field, m_data.data())); }
  void download() { if (m_ctx) FSGPU_CHECK(fsgpu_download(m_ctx, m_field, m_data.data())); }
};

namespace fsgpu_shim {
template <class F>
fsgpu_ctx* ctx_of(const F& f) {
  if (!f.bound()) check(1, "field is not bound to a device context (call init_grid(..., fs) first)");
  return f.context();
}
}  // namespace fsgpu_shim

// ---- host loops of src/ForEach.hpp:23-136 (for_each, for_each_i, for_each_a) ----------------------
// Drivers initialise fields and evaluate their test criteria with these; here they run serially on
// the HOST mirrors (Exec is accepted and ignored), so such code compiles unchanged.  Follow
// initialisation loops by fs.to_device() / field.upload(), precede evaluation loops by fs.to_host().
enum class Exec : std::uint8_t { Serial, Parallel, ParallelDynamic, ParallelGPU };
template <Index I_MIN, Index I_MAX, Exec EXEC = Exec::Serial, typename FUNC>
void for_each(FUNC&& f) {
  for (Index i = I_MIN; i < I_MAX; ++i) f(i);
}
template <Index I_MIN, Index I_MAX, Index J_MIN, Index J_MAX, Exec EXEC = Exec::Serial, typename FUNC>
void for_each(FUNC&& f) {
  for (Index i = I_MIN; i < I_MAX; ++i)
    for (Index j = J_MIN; j < J_MAX; ++j) f(i, j);
}
template <Exec EXEC = Exec::Serial, typename Float, Index N, Index NGHOST, typename FUNC>
void for_each_i(const Field1D<Float, N, NGHOST>&, FUNC&& f) { for_each<0, N, EXEC>(f); }
template <Exec EXEC = Exec::Serial, typename Float, Index N, Index NGHOST, typename FUNC>
void for_each_a(const Field1D<Float, N, NGHOST>&, FUNC&& f) { for_each<-NGHOST, N + NGHOST, EXEC>(f); }
template <Exec EXEC = Exec::Serial, typename Float, Index M, Index N, Index NGHOST, typename FUNC>
void for_each_i(const Field2D<Float, M, N, NGHOST>&, FUNC&& f) { for_each<0, M, 0, N, EXEC>(f); }
template <Exec EXEC = Exec::Serial, typename Float, Index M, Index N, Index NGHOST, typename FUNC>
void for_each_a(const Field2D<Float, M, N, NGHOST>&, FUNC&& f) {
  for_each<-NGHOST, M + NGHOST, -NGHOST, N + NGHOST, EXEC>(f);
}

// fill(field, value), src/Container.hpp:172-180: host mirror and device
template <typename Float, Index M, Index N, Index NGHOST>
void fill(Field2D<Float, M, N, NGHOST>& f, Float v) {
  for (Index k = 0; k < f.size(); ++k) f.get_data()[k] = v;
  f.upload();
}

// ---- State / FS (src/FS.hpp:15-52) ------------------------------------------------------------------
template <typename Float, Index NX, Index NY, Index NGHOST>
struct State {
  Field2D<Float, NX + 1, NY, NGHOST> rho_u_stag{};
  Field2D<Float, NX, NY + 1, NGHOST> rho_v_stag{};
  Field2D<Float, NX + 1, NY, NGHOST> U{};
  Field2D<Float, NX, NY + 1, NGHOST> V{};
};

template <typename Float, Index NX, Index NY, Index NGHOST>
struct FS {
  static_assert(NGHOST == 1, "the stepping path uses exactly one ghost cell (BoundaryConditions.hpp:219)");
  Float visc_gas{};
  Float visc_liquid{};
  Float rho_gas{};
  Float rho_liquid{};

  Field1D<Float, NX + 1, NGHOST> x{};
  Field1D<Float, NX, NGHOST> xm{};
  Float dx{};
  Field1D<Float, NY + 1, NGHOST> y{};
  Field1D<Float, NY, NGHOST> ym{};
  Float dy{};

  Field2D<Float, NX, NY, NGHOST> visc{};
  Field2D<Float, NX, NY, NGHOST> p{};
  Float sigma{};
  Field2D<Float, NX + 1, NY, NGHOST> p_jump_u_stag{};
  Field2D<Float, NX, NY + 1, NGHOST> p_jump_v_stag{};

  State<Float, NX, NY, NGHOST> old{};
  State<Float, NX, NY, NGHOST> curr{};

  // ---- device side (not in the reference) ----
  fsgpu_ctx* ctx = nullptr;
  int device     = 0;
  ~FS() { if (ctx) fsgpu_destroy(ctx); }

  // created by init_grid (dx, dy are known then)
  void create_context() {
    if (ctx) return;
    FSGPU_CHECK(fsgpu_create(NX, NY, dx, dy, device, &ctx));
    FSGPU_CHECK(fsgpu_set_props(ctx, rho_gas, rho_liquid, visc_gas, visc_liquid, sigma));
    visc.bind(ctx, FSGPU_VISC); p.bind(ctx, FSGPU_P);
    p_jump_u_stag.bind(ctx, FSGPU_P_JUMP_U); p_jump_v_stag.bind(ctx, FSGPU_P_JUMP_V);
    old.rho_u_stag.bind(ctx, FSGPU_RHO_U_OLD); old.rho_v_stag.bind(ctx, FSGPU_RHO_V_OLD);
    old.U.bind(ctx, FSGPU_U_OLD); old.V.bind(ctx, FSGPU_V_OLD);
    curr.rho_u_stag.bind(ctx, FSGPU_RHO_U); curr.rho_v_stag.bind(ctx, FSGPU_RHO_V);
    curr.U.bind(ctx, FSGPU_U); curr.V.bind(ctx, FSGPU_V);
  }
  // host mirrors -> HBM (after host-side initialisation) and back (before inspection / output)
  void to_device() {
    FSGPU_CHECK(fsgpu_set_props(ctx, rho_gas, rho_liquid, visc_gas, visc_liquid, sigma));
    visc.upload(); p.upload(); p_jump_u_stag.upload(); p_jump_v_stag.upload();
    old.rho_u_stag.upload(); old.rho_v_stag.upload(); old.U.upload(); old.V.upload();
    curr.rho_u_stag.upload(); curr.rho_v_stag.upload(); curr.U.upload(); curr.V.upload();
  }
  void to_host() {
    visc.download(); p.download();
    old.rho_u_stag.download(); old.rho_v_stag.download(); old.U.download(); old.V.download();
    curr.rho_u_stag.download(); curr.rho_v_stag.download(); curr.U.download(); curr.V.download();
  }
};

// init_grid, src/FS.hpp:55-73 (+ creation of the device context)
template <typename Float, Index NX, Index NY, Index NGHOST>
void init_grid(Float x_min, Float x_max, Index nx, Float y_min, Float y_max, Index ny,
               FS<Float, NX, NY, NGHOST>& fs) {
  fs.dx = (x_max - x_min) / static_cast<Float>(nx);
  fs.dy = (y_max - y_min) / static_cast<Float>(ny);
  for (Index i = -NGHOST; i < NX + 1 + NGHOST; ++i) fs.x(i) = x_min + static_cast<Float>(i) * fs.dx;
  for (Index j = -NGHOST; j < NY + 1 + NGHOST; ++j) fs.y(j) = y_min + static_cast<Float>(j) * fs.dy;
  for (Index i = -NGHOST; i < NX + NGHOST; ++i) fs.xm(i) = (fs.x(i) + fs.x(i + 1)) / 2;
  for (Index j = -NGHOST; j < NY + NGHOST; ++j) fs.ym(j) = (fs.y(j) + fs.y(j + 1)) / 2;
  fs.create_context();
}

// calc_rho(fs) / calc_visc(fs), src/FS.hpp:569-589 (constant properties, on the device)
template <typename Float, Index NX, Index NY, Index NGHOST>
void calc_rho(FS<Float, NX, NY, NGHOST>& fs) {
  for (auto* f : {&fs.curr.rho_u_stag, &fs.old.rho_u_stag}) fill(*f, fs.rho_gas);
  for (auto* f : {&fs.curr.rho_v_stag, &fs.old.rho_v_stag}) fill(*f, fs.rho_gas);
}
template <typename Float, Index NX, Index NY, Index NGHOST>
void calc_visc(FS<Float, NX, NY, NGHOST>& fs) { fill(fs.visc, fs.visc_gas); }

// calc_rho(vf, fs) / calc_visc(vf, fs), src/FS.hpp:592-650: one device pass computes the face
// densities and the viscosity (incl. their Neumann ghosts); it runs on the calc_rho call
template <typename Float, Index NX, Index NY, Index NGHOST>
void calc_rho(Field2D<Float, NX, NY, NGHOST>& vf, FS<Float, NX, NY, NGHOST>& fs) {
  vf.bind(fs.ctx, FSGPU_VF);
  vf.upload();
  FSGPU_CHECK(fsgpu_set_props(fs.ctx, fs.rho_gas, fs.rho_liquid, fs.visc_gas, fs.visc_liquid, fs.sigma));
  FSGPU_CHECK(fsgpu_calc_rho_visc_vof(fs.ctx));
}
template <typename Float, Index NX, Index NY, Index NGHOST>
void calc_visc(const Field2D<Float, NX, NY, NGHOST>& vf, FS<Float, NX, NY, NGHOST>&) {
  if (vf.field_id() != FSGPU_VF) fsgpu_shim::check(1, "calc_visc(vf, fs) must follow calc_rho(vf, fs)");
}

// ---- boundary conditions (src/BoundaryConditions.hpp:14-224) --------------------------------------
template <typename Float>
struct Dirichlet {
  std::variant<Float, Float (*)(Float, Float)> U{}, V{};
};
struct Neumann { bool clipped = false; };
struct Periodic {};
struct Symmetry {};
template <typename Float>
using BCond_t = std::variant<Dirichlet<Float>, Neumann, Periodic, Symmetry>;
template <typename Float>
struct FlowBConds { BCond_t<Float> left, right, bottom, top; };

namespace fsgpu_shim {
template <typename Float, Index NX, Index NY>
void set_side(FS<Float, NX, NY, 1>& fs, int side, const BCond_t<Float>& bc, Float t) {
  if (const auto* d = std::get_if<Dirichlet<Float>>(&bc)) {
    // function-pointer profiles are evaluated on the host at the ghost-inclusive coordinates the
    // reference uses (left/right: U at ym(j), V at y(j); bottom/top: U at x(i), V at xm(i))
    const bool lr = side == FSGPU_LEFT || side == FSGPU_RIGHT;
    std::vector<double> up, vp;
    double uval = 0.0, vval = 0.0;
    if (std::holds_alternative<Float>(d->U)) uval = std::get<0>(d->U);
    else {
      auto f = std::get<1>(d->U);
      if (lr) for (Index j = -1; j <= NY; ++j) up.push_back(f(fs.ym(j), t));
      else for (Index i = -1; i <= NX + 1; ++i) up.push_back(f(fs.x(i), t));
    }
    if (std::holds_alternative<Float>(d->V)) vval = std::get<0>(d->V);
    else {
      auto f = std::get<1>(d->V);
      if (lr) for (Index j = -1; j <= NY + 1; ++j) vp.push_back(f(fs.y(j), t));
      else for (Index i = -1; i <= NX; ++i) vp.push_back(f(fs.xm(i), t));
    }
    FSGPU_CHECK(fsgpu_set_bconds(fs.ctx, side, FSGPU_BC_DIRICHLET, uval, vval, up.empty() ? nullptr : up.data(),
                                 vp.empty() ? nullptr : vp.data(), 0));
  } else if (const auto* n = std::get_if<Neumann>(&bc)) {
    FSGPU_CHECK(fsgpu_set_bconds(fs.ctx, side, FSGPU_BC_NEUMANN, 0, 0, nullptr, nullptr, n->clipped ? 1 : 0));
  } else if (std::holds_alternative<Periodic>(bc)) {
    FSGPU_CHECK(fsgpu_set_bconds(fs.ctx, side, FSGPU_BC_PERIODIC, 0, 0, nullptr, nullptr, 0));
  } else {
    FSGPU_CHECK(fsgpu_set_bconds(fs.ctx, side, FSGPU_BC_SYMMETRY, 0, 0, nullptr, nullptr, 0));
  }
}
template <typename Float>
bool time_dependent(const FlowBConds<Float>& b) {
  for (const BCond_t<Float>* s : {&b.left, &b.right, &b.bottom, &b.top})
    if (const auto* d = std::get_if<Dirichlet<Float>>(s))
      if (!std::holds_alternative<Float>(d->U) || !std::holds_alternative<Float>(d->V)) return true;
  return false;
}
}  // namespace fsgpu_shim

// apply_velocity_bconds(fs, bconds, t), src/BoundaryConditions.hpp:215-224.  The side table is sent
// to the device on the first call and again whenever a function-pointer profile may depend on t.
template <typename Float, Index NX, Index NY>
void apply_velocity_bconds(FS<Float, NX, NY, 1>& fs, const FlowBConds<Float>& bconds, Float t = -1.0) {
  static const void* sent_for = nullptr;
  static const void* sent_bc  = nullptr;
  static double sent_t        = 0.0;
  if (sent_for != &fs || sent_bc != &bconds || (fsgpu_shim::time_dependent(bconds) && sent_t != t)) {
    fsgpu_shim::set_side(fs, FSGPU_LEFT, bconds.left, t);
    fsgpu_shim::set_side(fs, FSGPU_RIGHT, bconds.right, t);
    fsgpu_shim::set_side(fs, FSGPU_BOTTOM, bconds.bottom, t);
    fsgpu_shim::set_side(fs, FSGPU_TOP, bconds.top, t);
    sent_for = &fs;
    sent_bc  = &bconds;
    sent_t   = t;
  }
  FSGPU_CHECK(fsgpu_apply_velocity_bconds(fs.ctx));
}

// ---- the stepping path (src/FS.hpp, src/Operators.hpp, src/Utility.hpp) ------------------------------
// adjust_dt, src/FS.hpp:100-128
template <typename Float, Index NX, Index NY, Index NGHOST>
Float adjust_dt(const FS<Float, NX, NY, NGHOST>& fs, Float cfl_max, Float dt_max) {
  double dt = 0.0;
  FSGPU_CHECK(fsgpu_adjust_dt(fs.ctx, cfl_max, dt_max, &dt));
  return dt;
}
// save_old_state / save_old_velocity / save_old_density, src/FS.hpp:76-97
template <typename Float, Index NX, Index NY, Index NGHOST>
void save_old_state(const State<Float, NX, NY, NGHOST>& curr, State<Float, NX, NY, NGHOST>& old) {
  (void)old;
  FSGPU_CHECK(fsgpu_save_old_state(fsgpu_shim::ctx_of(curr.U)));
}
template <typename Float, Index NX, Index NY, Index NGHOST>
void save_old_velocity(const State<Float, NX, NY, NGHOST>& curr, State<Float, NX, NY, NGHOST>&) {
  FSGPU_CHECK(fsgpu_save_old_velocity(fsgpu_shim::ctx_of(curr.U)));
}
template <typename Float, Index NX, Index NY, Index NGHOST>
void save_old_density(const State<Float, NX, NY, NGHOST>& curr, State<Float, NX, NY, NGHOST>&) {
  FSGPU_CHECK(fsgpu_save_old_density(fsgpu_shim::ctx_of(curr.U)));
}
// calc_mid_time(current, old), src/Operators.hpp:44-49.  The drivers call it once for U and once for
// V (scaling/LidDrivenFlow.cpp:128-129); the device pass averages both, so it runs on the U call.
template <typename Float, Index M, Index N, Index NGHOST>
void calc_mid_time(Field2D<Float, M, N, NGHOST>& current, const Field2D<Float, M, N, NGHOST>& old) {
  (void)old;
  if (current.field_id() == FSGPU_U) FSGPU_CHECK(fsgpu_calc_mid_time(fsgpu_shim::ctx_of(current)));
  else if (current.field_id() != FSGPU_V) fsgpu_shim::check(1, "calc_mid_time on a field that is not curr.U / curr.V");
}
// calc_dmomdt(fs, dmomUdt, dmomVdt), src/FS.hpp:158-315
template <typename Float, Index NX, Index NY, Index NGHOST>
void calc_dmomdt(const FS<Float, NX, NY, NGHOST>& fs, Field2D<Float, NX + 1, NY, NGHOST>& dmomUdt,
                 Field2D<Float, NX, NY + 1, NGHOST>& dmomVdt) {
  dmomUdt.bind(fs.ctx, FSGPU_DMOM_U);
  dmomVdt.bind(fs.ctx, FSGPU_DMOM_V);
  FSGPU_CHECK(fsgpu_calc_dmomdt(fs.ctx));
}
// update_velocity(dmomUdt, dmomVdt, dt, fs), src/FS.hpp:423-436
template <typename Float, Index NX, Index NY, Index NGHOST>
void update_velocity(const Field2D<Float, NX + 1, NY, NGHOST>&, const Field2D<Float, NX, NY + 1, NGHOST>&,
                     Float dt, FS<Float, NX, NY, NGHOST>& fs) {
  FSGPU_CHECK(fsgpu_update_velocity(fs.ctx, dt));
}
// calc_divergence(U, V, dx, dy, div), src/Operators.hpp:32-41
template <typename Float, Index NX, Index NY, Index NGHOST>
void calc_divergence(const Field2D<Float, NX + 1, NY, NGHOST>& U, const Field2D<Float, NX, NY + 1, NGHOST>& V,
                     Float, Float, Field2D<Float, NX, NY, NGHOST>& div) {
  if (U.field_id() != FSGPU_U || V.field_id() != FSGPU_V) fsgpu_shim::check(1, "calc_divergence expects fs.curr.U, fs.curr.V");
  div.bind(fsgpu_shim::ctx_of(U), FSGPU_DIV);
  FSGPU_CHECK(fsgpu_calc_divergence(fsgpu_shim::ctx_of(U)));
}
// shift_pressure_to_zero(dx, dy, delta_p), src/Operators.hpp:78-82
template <typename Float, Index NX, Index NY, Index NGHOST>
void shift_pressure_to_zero(Float, Float, Field2D<Float, NX, NY, NGHOST>& dp) {
  FSGPU_CHECK(fsgpu_shift_pressure_to_zero(fsgpu_shim::ctx_of(dp)));
}
// the projection every driver inlines: p += delta_p; U, V -= grad(delta_p) dt / rho
// (scaling/LidDrivenFlow.cpp:153-160; variable density: test/LaminarChannel.cpp:222-233)
template <typename Float, Index NX, Index NY, Index NGHOST>
void correct_velocity(FS<Float, NX, NY, NGHOST>& fs, const Field2D<Float, NX, NY, NGHOST>&, Float dt) {
  FSGPU_CHECK(fsgpu_correct_velocity(fs.ctx, dt));
}
// abs_max<INCLUDE_GHOST = false>(field), src/Utility.hpp:44-88: interior entries unless INCLUDE_GHOST
template <bool INCLUDE_GHOST = false, typename Float, Index M, Index N, Index NGHOST>
Float abs_max(const Field2D<Float, M, N, NGHOST>& f) {
  double v = 0.0;
  if (INCLUDE_GHOST) FSGPU_CHECK(fsgpu_abs_max_all(fsgpu_shim::ctx_of(f), f.field_id(), &v));
  else FSGPU_CHECK(fsgpu_abs_max(fsgpu_shim::ctx_of(f), f.field_id(), &v));
  return v;
}

// ---- rows either side of the single-phase path -------------------------------------
// calc_drhodt(fs, drho_u_stagdt, drho_v_stagdt), src/FS.hpp:318-406
template <typename Float, Index NX, Index NY, Index NGHOST>
void calc_drhodt(const FS<Float, NX, NY, NGHOST>& fs, Field2D<Float, NX + 1, NY, NGHOST>& drho_u_stagdt,
                 Field2D<Float, NX, NY + 1, NGHOST>& drho_v_stagdt) {
  drho_u_stagdt.bind(fs.ctx, FSGPU_DRHO_U);
  drho_v_stagdt.bind(fs.ctx, FSGPU_DRHO_V);
  FSGPU_CHECK(fsgpu_calc_drhodt(fs.ctx));
}
// update_density(drho_u_stagdt, drho_v_stagdt, dt, fs), src/FS.hpp:409-420
template <typename Float, Index NX, Index NY, Index NGHOST>
void update_density(const Field2D<Float, NX + 1, NY, NGHOST>&, const Field2D<Float, NX, NY + 1, NGHOST>&,
                    Float dt, FS<Float, NX, NY, NGHOST>& fs) {
  FSGPU_CHECK(fsgpu_update_density(fs.ctx, dt));
}
// apply_neumann_bconds(field), src/BoundaryConditions.hpp:227-252
template <typename Float, Index M, Index N, Index NGHOST>
void apply_neumann_bconds(Field2D<Float, M, N, NGHOST>& f) {
  FSGPU_CHECK(fsgpu_apply_neumann_bconds(fsgpu_shim::ctx_of(f), f.field_id()));
}
// calc_pressure_jump(vf, curv, interface_length, fs), src/FS.hpp:439-466.  The three cell fields
// come from the VOF module (out of scope here): bind_cell_input uploads a host-filled field.
template <typename Float, Index NX, Index NY, Index NGHOST>
void bind_cell_input(FS<Float, NX, NY, NGHOST>& fs, Field2D<Float, NX, NY, NGHOST>& f, int field) {
  f.bind(fs.ctx, field);
  f.upload();
}
template <typename Float, Index NX, Index NY, Index NGHOST>
void calc_pressure_jump(Field2D<Float, NX, NY, NGHOST>& vf, Field2D<Float, NX, NY, NGHOST>& curv,
                        Field2D<Float, NX, NY, NGHOST>& interface_length, FS<Float, NX, NY, NGHOST>& fs) {
  if (!vf.bound()) bind_cell_input(fs, vf, FSGPU_VF);
  if (!curv.bound()) bind_cell_input(fs, curv, FSGPU_CURV);
  if (!interface_length.bound()) bind_cell_input(fs, interface_length, FSGPU_INTERFACE_LENGTH);
  FSGPU_CHECK(fsgpu_calc_pressure_jump(fs.ctx));
}
// the capillary block the two-phase drivers inline (test/StationaryDrop.cpp:331-353)
template <typename Float, Index NX, Index NY, Index NGHOST>
void add_capillary_divergence(FS<Float, NX, NY, NGHOST>& fs, Field2D<Float, NX, NY, NGHOST>& div, Float dt) {
  if (div.field_id() != FSGPU_DIV) fsgpu_shim::check(1, "add_capillary_divergence expects the field calc_divergence wrote");
  FSGPU_CHECK(fsgpu_capillary_divergence(fs.ctx, dt));
}
// correct_velocity_ib_implicit_euler / update_velocity_ib_semi_analytical, src/IB.hpp:110-186
template <typename Float, Index NX, Index NY, Index NGHOST>
void correct_velocity_ib_implicit_euler(Field2D<Float, NX + 1, NY, NGHOST>& ib_corr_u_stag,
                                        Field2D<Float, NX, NY + 1, NGHOST>& ib_corr_v_stag, Float dt,
                                        FS<Float, NX, NY, NGHOST>& fs) {
  if (!ib_corr_u_stag.bound()) { ib_corr_u_stag.bind(fs.ctx, FSGPU_IB_CORR_U); ib_corr_u_stag.upload(); }
  if (!ib_corr_v_stag.bound()) { ib_corr_v_stag.bind(fs.ctx, FSGPU_IB_CORR_V); ib_corr_v_stag.upload(); }
  FSGPU_CHECK(fsgpu_correct_velocity_ib_implicit_euler(fs.ctx, dt));
}
template <typename Float, Index NX, Index NY, Index NGHOST>
void update_velocity_ib_semi_analytical(const Field2D<Float, NX + 1, NY, NGHOST>&,
                                        const Field2D<Float, NX, NY + 1, NGHOST>&, Float dt,
                                        Field2D<Float, NX + 1, NY, NGHOST>& ib_corr_u_stag,
                                        Field2D<Float, NX, NY + 1, NGHOST>& ib_corr_v_stag,
                                        FS<Float, NX, NY, NGHOST>& fs) {
  if (!ib_corr_u_stag.bound()) { ib_corr_u_stag.bind(fs.ctx, FSGPU_IB_CORR_U); ib_corr_u_stag.upload(); }
  if (!ib_corr_v_stag.bound()) { ib_corr_v_stag.bind(fs.ctx, FSGPU_IB_CORR_V); ib_corr_v_stag.upload(); }
  FSGPU_CHECK(fsgpu_update_velocity_ib_semi_analytical(fs.ctx, dt));
}
// interpolate_U / interpolate_V / interpolate_UV_staggered_field, src/Operators.hpp:10-29: one device
// pass fills all three cell-centred fields; it runs on the interpolate_U call
template <typename Float, Index NX, Index NY, Index NGHOST>
void interpolate_U(const Field2D<Float, NX + 1, NY, NGHOST>& U, Field2D<Float, NX, NY, NGHOST>& Ui) {
  Ui.bind(fsgpu_shim::ctx_of(U), FSGPU_UI);
  FSGPU_CHECK(fsgpu_interpolate_cell_centres(fsgpu_shim::ctx_of(U)));
}
template <typename Float, Index NX, Index NY, Index NGHOST>
void interpolate_V(const Field2D<Float, NX, NY + 1, NGHOST>& V, Field2D<Float, NX, NY, NGHOST>& Vi) {
  Vi.bind(fsgpu_shim::ctx_of(V), FSGPU_VI);
}
template <typename Float, Index NX, Index NY, Index NGHOST>
void interpolate_UV_staggered_field(const Field2D<Float, NX + 1, NY, NGHOST>& u_stag,
                                    const Field2D<Float, NX, NY + 1, NGHOST>&,
                                    Field2D<Float, NX, NY, NGHOST>& interp) {
  if (u_stag.field_id() != FSGPU_RHO_U) fsgpu_shim::check(1, "interpolate_UV_staggered_field expects curr.rho_u_stag / rho_v_stag");
  interp.bind(fsgpu_shim::ctx_of(u_stag), FSGPU_RHOI);
}
// calc_conserved_quantities(fs, mass, momentum_x, momentum_y), src/FS.hpp:653-676
template <typename Float, Index NX, Index NY, Index NGHOST>
void calc_conserved_quantities(const FS<Float, NX, NY, NGHOST>& fs, Float& mass, Float& momentum_x,
                               Float& momentum_y) {
  FSGPU_CHECK(fsgpu_calc_conserved_quantities(fs.ctx, &mass, &momentum_x, &momentum_y));
}

// eval_flow_field_at / bilinear_interpolate, src/Operators.hpp:85-161, on device-bound cell fields: the reference's
// one-point signatures (one kernel per call) and batched forms for callers that sample many points per step
template <typename Float, Index NX, Index NY, Index NGHOST>
void eval_flow_field_at(const Field1D<Float, NX, NGHOST>& xm, const Field1D<Float, NY, NGHOST>& ym,
                        const Field2D<Float, NX, NY, NGHOST>& Ui, const Field2D<Float, NX, NY, NGHOST>& Vi, Index n,
                        const Float* x, const Float* y, Float* u, Float* v) {
  if (Ui.field_id() != FSGPU_UI || Vi.field_id() != FSGPU_VI) fsgpu_shim::check(1, "eval_flow_field_at expects the fields interpolate_U / interpolate_V filled");
  FSGPU_CHECK(fsgpu_eval_flow_field_at(fsgpu_shim::ctx_of(Ui), xm.get_data(), ym.get_data(), n, x, y, u, v));
}
template <typename Float, Index NX, Index NY, Index NGHOST>
[[nodiscard]] std::pair<Float, Float> eval_flow_field_at(const Field1D<Float, NX, NGHOST>& xm, const Field1D<Float, NY, NGHOST>& ym,
                                                         const Field2D<Float, NX, NY, NGHOST>& Ui,
                                                         const Field2D<Float, NX, NY, NGHOST>& Vi, Float x, Float y) {
  Float u = 0, v = 0;
  eval_flow_field_at(xm, ym, Ui, Vi, Index{1}, &x, &y, &u, &v);
  return {u, v};
}
template <typename Float, Index NX, Index NY, Index NGHOST>
[[nodiscard]] Float bilinear_interpolate(const Field1D<Float, NX, NGHOST>& xm, const Field1D<Float, NY, NGHOST>& ym,
                                         const Field2D<Float, NX, NY, NGHOST>& field, Float x, Float y) {
  Float out = 0;
  FSGPU_CHECK(fsgpu_bilinear_interpolate(fsgpu_shim::ctx_of(field), field.field_id(), xm.get_data(), ym.get_data(), 1, &x,
                                         &y, &out));
  return out;
}

// The channel drivers' corrections of the two ghost columns U(-1, .) and U(NX+1, .) after
// apply_velocity_bconds.  The reference drivers inline them as host loops (calc_inflow_outflow + for_each_a over
// fs.ym); with the state in HBM they are these two calls.  inflow / outflow are the fluxes BEFORE the correction.
//   correct_outflow(fs, false, ...)  test/LaminarChannel.cpp:55-66,201-205
//   correct_outflow(fs, true, ...)   test/CustomDirichletChannel.cpp:61-72,197-202, scaling/Channel.cpp:208-213
//   force_total_flow(fs, Q, ...)     test/PeriodicChannel.cpp:56-67,193-203, scaling/Channel.cpp:196-206
template <typename Float, Index NX, Index NY, Index NGHOST>
void correct_outflow(FS<Float, NX, NY, NGHOST>& fs, bool dy_weighted, Float& inflow, Float& outflow, Float& mass_error) {
  FSGPU_CHECK(fsgpu_flow_control(fs.ctx, dy_weighted ? FSGPU_FLOW_OUTFLOW_FIX_DY : FSGPU_FLOW_OUTFLOW_FIX, 0.0,
                                 &inflow, &outflow));
  mass_error = outflow - inflow;
}
template <typename Float, Index NX, Index NY, Index NGHOST>
void force_total_flow(FS<Float, NX, NY, NGHOST>& fs, Float total_flow, Float& inflow, Float& outflow) {
  FSGPU_CHECK(fsgpu_flow_control(fs.ctx, FSGPU_FLOW_FORCE_TOTAL, total_flow, &inflow, &outflow));
}

// has_nan / has_inf / has_nan_or_inf, src/Container.hpp:191-204, on a device-bound field (one reduction pass,
// no copy to the host)
namespace fsgpu_shim {
template <class F>
void count_nonfinite(const F& f, std::int64_t& n_nan, std::int64_t& n_inf) {
  FSGPU_CHECK(fsgpu_count_nonfinite(ctx_of(f), f.field_id(), &n_nan, &n_inf));
}
}  // namespace fsgpu_shim
template <typename Float, Index M, Index N, Index NGHOST>
bool has_nan(const Field2D<Float, M, N, NGHOST>& f) {
  std::int64_t a = 0, b = 0;
  fsgpu_shim::count_nonfinite(f, a, b);
  return a > 0;
}
template <typename Float, Index M, Index N, Index NGHOST>
bool has_inf(const Field2D<Float, M, N, NGHOST>& f) {
  std::int64_t a = 0, b = 0;
  fsgpu_shim::count_nonfinite(f, a, b);
  return b > 0;
}
template <typename Float, Index M, Index N, Index NGHOST>
bool has_nan_or_inf(const Field2D<Float, M, N, NGHOST>& f) {
  std::int64_t a = 0, b = 0;
  fsgpu_shim::count_nonfinite(f, a, b);
  return a + b > 0;
}

// integrate / L1_norm (src/Operators.hpp:51-76) and calc_grad_of_centered_points (:164-201) on device-bound
// fields; the gradient's outputs are bound to FSGPU_GRAD_X / FSGPU_GRAD_Y unless already bound elsewhere
template <bool INCLUDE_GHOST = false, typename Float, Index M, Index N, Index NGHOST>
Float integrate(Float, Float, const Field2D<Float, M, N, NGHOST>& field) {
  double out = 0.0;
  FSGPU_CHECK(fsgpu_integrate(fsgpu_shim::ctx_of(field), field.field_id(), INCLUDE_GHOST ? 1 : 0, &out));
  return out;
}
template <bool INCLUDE_GHOST = false, typename Float, Index M, Index N, Index NGHOST>
Float L1_norm(Float, Float, const Field2D<Float, M, N, NGHOST>& field) {
  double out = 0.0;
  FSGPU_CHECK(fsgpu_l1_norm(fsgpu_shim::ctx_of(field), field.field_id(), INCLUDE_GHOST ? 1 : 0, &out));
  return out;
}
template <typename Float, Index NX, Index NY, Index NGHOST>
void calc_grad_of_centered_points(const Field2D<Float, NX, NY, NGHOST>& f, Float, Float,
                                  Field2D<Float, NX, NY, NGHOST>& dfdx, Field2D<Float, NX, NY, NGHOST>& dfdy) {
  fsgpu_ctx* ctx = fsgpu_shim::ctx_of(f);
  if (!dfdx.bound()) dfdx.bind(ctx, f.field_id() == FSGPU_GRAD_X ? FSGPU_GRAD_XX : FSGPU_GRAD_X);
  if (!dfdy.bound()) dfdy.bind(ctx, f.field_id() == FSGPU_GRAD_Y ? FSGPU_GRAD_YY : FSGPU_GRAD_Y);
  FSGPU_CHECK(fsgpu_calc_grad_of_centered_points(ctx, f.field_id(), dfdx.field_id(), dfdy.field_id()));
}

// ---- output (src/IO.hpp:178-269, src/VTKWriter.hpp) -----------------------------------------------
// The files are written from the device buffers and are byte-identical to the reference's.
// to_npy(filename, Field1D / Field2D): a field bound to a device field is written from HBM, any other
// from its host array.
template <typename Float, Index N, Index NGHOST>
[[nodiscard]] bool to_npy(const std::string& filename, const Field1D<Float, N, NGHOST>& v) {
  return fsgpu_npy_write_host(filename.c_str(), v.get_data(), v.size(), 0) == 0;
}
template <typename Float, Index M, Index N, Index NGHOST>
[[nodiscard]] bool to_npy(const std::string& filename, const Field2D<Float, M, N, NGHOST>& f) {
  if (f.bound()) return fsgpu_to_npy(f.context(), f.field_id(), filename.c_str()) == 0;
  return fsgpu_npy_write_host(filename.c_str(), f.get_data(), M + 2 * NGHOST, N + 2 * NGHOST) == 0;
}
// to_npy(directory, fs), src/IO.hpp:231-269 (the directory is created, not emptied first)
template <typename Float, Index NX, Index NY, Index NGHOST>
[[nodiscard]] bool to_npy(const std::string& directory, const FS<Float, NX, NY, NGHOST>& fs) {
  return fsgpu_to_npy_state(fs.ctx, directory.c_str(), fs.x.get_data(), fs.y.get_data()) == 0;
}

// VTKWriter (src/VTKWriter.hpp:14-152), the drivers' DataWriter when HDF5 is off (src/IO.hpp:13-16).
// Registered fields must be device-bound cell fields: fs.p, fs.visc, and the div / Ui / Vi / rhoi / vf /
// curv scratch fields once they went through calc_divergence / interpolate_* / calc_rho(vf, fs) /
// calc_pressure_jump.  The file counter is per writer (the reference's is per template instantiation).
template <typename Float, Index NX, Index NY, Index NGHOST>
class VTKWriter {
  using Cell = Field2D<Float, NX, NY, NGHOST>;
  std::string m_output_dir;
  const Field1D<Float, NX + 1, NGHOST>* m_x;
  const Field1D<Float, NY + 1, NGHOST>* m_y;
  std::vector<std::string> m_scalar_names, m_vector_names;
  std::vector<const Cell*> m_scalar_values;
  std::vector<std::array<const Cell*, 2>> m_vector_values;
  Index m_write_counter = 0;

 public:
  VTKWriter(std::string output_dir, const Field1D<Float, NX + 1, NGHOST>* x, const Field1D<Float, NY + 1, NGHOST>* y)
      : m_output_dir(std::move(output_dir)), m_x(x), m_y(y) {}
  VTKWriter(const VTKWriter&)            = delete;
  VTKWriter& operator=(const VTKWriter&) = delete;
  void add_scalar(std::string name, const Cell* value) {
    m_scalar_names.emplace_back(std::move(name));
    m_scalar_values.push_back(value);
  }
  void add_vector(std::string name, const Cell* x_value, const Cell* y_value) {
    m_vector_names.emplace_back(std::move(name));
    m_vector_values.push_back({x_value, y_value});
  }
  bool write(Float t = -1.0) {
    char name[32];
    std::snprintf(name, sizeof name, "/state_%06lld.vtk", static_cast<long long>(m_write_counter++));
    fsgpu_ctx* ctx = nullptr;
    std::vector<int> sf, vf;
    std::vector<const char*> sn, vn;
    auto use = [&](const Cell* f, std::vector<int>& ids) {
      ctx = fsgpu_shim::ctx_of(*f);
      ids.push_back(f->field_id());
    };
    for (size_t k = 0; k < m_scalar_values.size(); ++k) { use(m_scalar_values[k], sf); sn.push_back(m_scalar_names[k].c_str()); }
    for (size_t k = 0; k < m_vector_values.size(); ++k) {
      use(m_vector_values[k][0], vf); use(m_vector_values[k][1], vf);
      vn.push_back(m_vector_names[k].c_str());
    }
    if (!ctx) fsgpu_shim::check(1, "VTKWriter::write: no field registered");
    return fsgpu_write_vtk(ctx, (m_output_dir + name).c_str(), t, m_x->get_data(), m_y->get_data(),
                           static_cast<int>(sf.size()), sf.data(), sn.data(), static_cast<int>(vn.size()),
                           vf.data(), vn.data()) == 0;
  }
};
template <typename Float, Index NX, Index NY, Index NGHOST>
using DataWriter = VTKWriter<Float, NX, NY, NGHOST>;

// ---- host-side driver utilities: Monitor (src/Monitor.hpp), should_save / get_output_directory /
// ---- init_output_directory (src/IO.hpp:97-175).  Plain host code, same observable behaviour: the
// ---- monitor log is byte-identical to the reference's (tests/test_output_cpu.py).
template <typename Float>
class Monitor {
  struct Column {
    std::string name;
    const Float* f;
    const Index* i;
    size_t width;  // max(len(name), 13): 13 characters hold a `.6e` number
  };
  std::FILE* m_file = nullptr;
  std::vector<Column> m_cols;
  bool m_header_done = false;

  // text centred in `width` characters; an odd remainder goes to the right (as std::format's '^')
  static std::string centred(const std::string& text, size_t width) {
    if (text.size() >= width) return text;
    const size_t pad = width - text.size(), left = pad / 2;
    return std::string(left, ' ') + text + std::string(pad - left, ' ');
  }

 public:
  explicit Monitor(const std::string& filename) : m_file(std::fopen(filename.c_str(), "w")) {
    if (!m_file) {
      std::fprintf(stderr, "Could not open monitor file `%s`\n", filename.c_str());
      std::exit(1);
    }
  }
  Monitor(const Monitor&)            = delete;
  Monitor& operator=(const Monitor&) = delete;
  ~Monitor() { if (m_file) std::fclose(m_file); }

  void add_variable(const Float* variable, std::string name) {
    const size_t w = name.size() > 13 ? name.size() : 13;
    m_cols.push_back(Column{std::move(name), variable, nullptr, w});
  }
  void add_variable(const Index* variable, std::string name) {
    const size_t w = name.size() > 13 ? name.size() : 13;
    m_cols.push_back(Column{std::move(name), nullptr, variable, w});
  }
  void write() {
    if (m_cols.empty()) return;
    std::string line;
    if (!m_header_done) {
      line = "| ";
      for (const Column& c : m_cols) line += centred(c.name, c.width) + " | ";
      line += "\n|";
      for (const Column& c : m_cols) line += std::string(c.width + 2, '-') + "|";
      line += "\n";
      m_header_done = true;
    }
    line += "| ";
    char num[64];
    for (const Column& c : m_cols) {
      if (c.f) std::snprintf(num, sizeof num, "%.6e", static_cast<double>(*c.f));
      else std::snprintf(num, sizeof num, "%lld", static_cast<long long>(*c.i));
      line += centred(num, c.width) + " | ";
    }
    line += "\n";
    std::fwrite(line.data(), 1, line.size(), m_file);
    std::fflush(m_file);
  }
};

// should_save(t, dt, dt_write, t_end), src/IO.hpp:97-108: true when a multiple of dt_write falls into the step
// that just ended, and once at t_end (the last save time is remembered across calls, as in the reference)
template <typename Float>
[[nodiscard]] bool should_save(Float t, Float dt, Float dt_write, Float t_end) {
  constexpr Float SAFE = 1e-6;
  static Float last    = -1.0;
  const bool at_end    = std::abs(t - t_end) < SAFE;
  const bool crossed   = std::fmod(t + SAFE * dt, dt_write) < dt * (1.0 - SAFE);
  if (!(crossed || at_end)) return false;
  if (at_end && std::abs(t - last) < SAFE) return false;
  last = t;
  return true;
}

#ifndef FS_BASE_DIR
#define FS_BASE_DIR "."
#endif
// "<base_dir>/<subdir>/<source file name without directory and extension>/", src/IO.hpp:114-155
[[nodiscard]] inline std::string get_output_directory(std::string_view subdir = "output", std::string_view base_dir = FS_BASE_DIR,
                                                      std::source_location loc = std::source_location::current()) {
  const std::string_view file = loc.file_name();
  const size_t slash = file.rfind('/');
  std::string_view stem = slash == std::string_view::npos ? file : file.substr(slash + 1);
  const size_t dot = stem.rfind('.');
  // the reference drops one character more than the extension when there is no '.', keep real names intact
  if (dot != std::string_view::npos) stem = stem.substr(0, dot);
  return std::string(base_dir) + "/" + std::string(subdir) + "/" + std::string(stem) + "/";
}
// init_output_directory, src/IO.hpp:158-175: the directory is REMOVED with everything in it and created anew
[[nodiscard]] inline bool init_output_directory(std::string_view directory_name) {
  std::error_code ec;
  std::filesystem::remove_all(directory_name, ec);
  if (ec) return false;
  std::filesystem::create_directories(directory_name, ec);
  return !ec;
}

// ---- LinearSolver_StructHypre (src/LinearSolver_StructHypre.hpp:28-192) ----------------------------
enum class HypreSolver { GMRES, PCG, BiCGSTAB };   // only PCG + PFMG (the reference's defaults) exist here
enum class HyprePrecond { SMG, PFMG, NONE };
enum class PSDirichlet { NONE, LEFT, RIGHT, BOTTOM, TOP };  // src/LinearSystem.hpp:10

template <typename Float, Index NX, Index NY, Index NGHOST>
class LinearSolver_StructHypre {
  fsgpu_ps* m_ps   = nullptr;
  fsgpu_ctx* m_ctx = nullptr;
  fsgpu_ps_opts m_opts{};
  void create(fsgpu_ctx* ctx) {
    if (!m_ps) FSGPU_CHECK(fsgpu_ps_create(ctx, &m_opts, &m_ps));
    m_ctx = ctx;
  }
 public:
  LinearSolver_StructHypre(Float tol, int max_iter, HypreSolver solver = HypreSolver::PCG,
                           HyprePrecond precond = HyprePrecond::PFMG, PSDirichlet psd = PSDirichlet::NONE) {
    if (solver != HypreSolver::PCG || precond != HyprePrecond::PFMG)
      fsgpu_shim::check(1, "LinearSolver_StructHypre: only PCG + PFMG is implemented on the GPU");
    m_opts.tol = tol; m_opts.max_iter = max_iter; m_opts.psd = static_cast<int>(psd);
    m_opts.relax = 0; m_opts.max_levels = 0; m_opts.skip_relax = -1;
  }
  LinearSolver_StructHypre(const LinearSolver_StructHypre&)            = delete;
  LinearSolver_StructHypre& operator=(const LinearSolver_StructHypre&) = delete;
  ~LinearSolver_StructHypre() { if (m_ps) fsgpu_ps_destroy(m_ps); }

  void set_pressure_operator(const FS<Float, NX, NY, NGHOST>& fs) {   // :91-106
    create(fs.ctx);
    FSGPU_CHECK(fsgpu_ps_set_operator(m_ps));
  }
  void set_pressure_rhs(const FS<Float, NX, NY, NGHOST>& fs, const Field2D<Float, NX, NY, NGHOST>& div,
                        Float dt) {                                     // :109-116
    if (div.field_id() != FSGPU_DIV) fsgpu_shim::check(1, "set_pressure_rhs expects the field calc_divergence wrote");
    create(fs.ctx);
    FSGPU_CHECK(fsgpu_ps_set_rhs(m_ps, dt));
  }
  // :119-192.  Non-convergence only warns, as in the reference (:177-183)
  bool solve(Field2D<Float, NX, NY, NGHOST>& sol, Float* residual = nullptr, Index* num_iter = nullptr) {
    double res = 0.0;
    int it     = 0;
    if (!m_ps) fsgpu_shim::check(1, "solve before set_pressure_operator");
    FSGPU_CHECK(fsgpu_ps_solve(m_ps, &res, &it));
    sol.bind(m_ctx, FSGPU_DELTA_P);
    if (residual) *residual = res;
    if (num_iter) *num_iter = it;
#ifndef FS_SILENCE_CONV_WARN
    if (it >= m_opts.max_iter && res >= m_opts.tol)
      std::fprintf(stderr, "[WARN] pressure solve did not converge: residual %.6e after %d iterations\n", res, it);
#endif
    return true;
  }
  fsgpu_ps* handle() { return m_ps; }
};

#endif  // FSGPU_SHIM_FLUIDSOLVERGPU_HPP_
