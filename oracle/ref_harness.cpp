// oracle/ref_harness.cpp -- TEST INFRASTRUCTURE ONLY.
//
// extern "C" wrapper around the UNMODIFIED reference (headers under /root/reference/src and the
// vendored HYPRE 3.0.0), compiled by oracle/Makefile into oracle/_ref/libfsref.so.  No reference
// source is copied: this file only #includes the reference headers where they lie and calls their
// public functions.  The reference fixes grid sizes at compile time (template parameters), so the
// sizes in ref_harness_sizes.inc are instantiated and dispatched on at run time.
//
// Used by: tests/ (to pin oracle/fs_oracle.c and to generate tests/golden/), bench.py's
// cpu_baseline leg and `bench.py --impl reference`.  Never by the product path.
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cmath>
#include <cstring>
#include <vector>
#include <memory>

#include <omp.h>

#include "FS.hpp"
#include "IB.hpp"
#include "IO.hpp"
#include "LinearSolver_StructHypre.hpp"
#include "Monitor.hpp"
#include "Operators.hpp"
#include "Quadrature.hpp"
#include "Utility.hpp"

#include "fs_oracle.h"

namespace {

using Float = double;

// The reference's Dirichlet condition takes plain function pointers `Float (*)(Float coord, Float t)`
// (src/BoundaryConditions.hpp:15-17).  The harness receives per-point tables (fso_bconds::u_prof /
// v_prof, evaluated by the caller at the ghost-inclusive coordinates of the side), so each
// (side, component) gets one static function that looks its table up by coordinate.
struct ProfTable {
  const double* v = nullptr;
  double c0 = 0.0, h = 1.0;  // coordinate of entry 0 and spacing
  long n = 0;
};
ProfTable g_prof[4][2];
template <int SIDE, int COMP>
Float prof_fn(Float coord, Float /*t*/) {
  const ProfTable& p = g_prof[SIDE][COMP];
  long k = std::lround((coord - p.c0) / p.h);
  if (k < 0) k = 0;
  if (k >= p.n) k = p.n - 1;
  return p.v[k];
}
using ProfFn = Float (*)(Float, Float);
constexpr ProfFn PROF_FNS[4][2] = {{&prof_fn<0, 0>, &prof_fn<0, 1>}, {&prof_fn<1, 0>, &prof_fn<1, 1>},
                                   {&prof_fn<2, 0>, &prof_fn<2, 1>}, {&prof_fn<3, 0>, &prof_fn<3, 1>}};

struct SessionBase {
  virtual ~SessionBase()                                                       = default;
  virtual void load(const fso_fs* in)                                          = 0;
  virtual void store(fso_fs* out) const                                        = 0;
  virtual void set_bconds(const fso_bconds* bc)                                = 0;
  virtual double adjust_dt(double cfl, double dt_max)                          = 0;
  virtual void save_old_state()                                                = 0;
  virtual void calc_mid_time()                                                 = 0;
  virtual void calc_dmomdt(double* dU, double* dV)                             = 0;
  virtual void update_velocity(const double* dU, const double* dV, double dt)  = 0;
  virtual void apply_bconds()                                                  = 0;
  virtual void calc_divergence(double* div)                                    = 0;
  virtual void fill_operator(int psd, double* op)                              = 0;
  virtual void fill_rhs(const double* div, double dt, int psd, double* rhs)    = 0;
  virtual void shift_pressure(double* dp)                                      = 0;
  virtual void calc_drhodt(double* dru, double* drv)                           = 0;
  virtual void rho_visc_vof(const double* vf)                                  = 0;
  virtual void solver_create(double tol, int max_iter, int psd)                = 0;
  virtual void solver_set_operator()                                           = 0;
  virtual int solver_solve(const double* div, double dt, double* dp, double* res) = 0;
  virtual int step(const fso_step_opts* o, double dt, double* res, double* phase_secs) = 0;
  virtual void diagnostics(double* umax, double* vmax, double* divmax)         = 0;
  virtual void update_density(const double* dru, const double* drv, double dt) = 0;
  virtual void pressure_jump(const double* vf, const double* curv, const double* ilen) = 0;
  virtual void interpolate_uv(double* ui, double* vi)                          = 0;
  virtual void conserved(double* mass, double* mx, double* my)                 = 0;
  virtual double flow_control(int mode, double total_flow, double* io)         = 0;
  virtual int to_npy(const char* dir)                                          = 0;
  virtual void ib_channel_correction(double y0, double y1, double* ibu, double* ibv) = 0;
  virtual void vf_drop(double* vf)                                             = 0;
  virtual void sample(int mode, int n, const double* px, const double* py, double* a, double* b) = 0;
  virtual double integrate_cell(const double* f, int include_ghost, int l1)    = 0;
  virtual void grad_centered(const double* f, double* dfdx, double* dfdy)      = 0;
  virtual int write_vtk(const char* dir, double t)                             = 0;
  virtual void ib_implicit_euler(const double* ibu, const double* ibv, double dt) = 0;
  virtual void ib_semi_analytical(const double* dU, const double* dV, double dt, const double* ibu,
                                  const double* ibv)                           = 0;
};

template <Index NX, Index NY>
struct Session final : SessionBase {
  static constexpr Index G = 1;
  FS<Float, NX, NY, G> fs{};
  Field2D<Float, NX + 1, NY, G> dU{};
  Field2D<Float, NX, NY + 1, G> dV{};
  Field2D<Float, NX, NY, G> div{};
  Field2D<Float, NX, NY, G> dp{};
  FlowBConds<Float> bconds{Neumann{}, Neumann{}, Neumann{}, Neumann{}};
  std::unique_ptr<LinearSolver_StructHypre<Float, NX, NY, G>> ps{};
  int psd_ = 0;

  template <typename F>
  static void cp_in(F& f, const double* src) {
    if (src) std::memcpy(f.get_data(), src, sizeof(double) * static_cast<size_t>(f.size()));
  }
  template <typename F>
  static void cp_out(const F& f, double* dst) {
    if (dst) std::memcpy(dst, f.get_data(), sizeof(double) * static_cast<size_t>(f.size()));
  }

  void load(const fso_fs* in) override {
    fs.visc_gas = in->visc_gas; fs.visc_liquid = in->visc_liquid;
    fs.rho_gas = in->rho_gas; fs.rho_liquid = in->rho_liquid; fs.sigma = in->sigma;
    // init_grid (src/FS.hpp:55-73) on [0, nx*dx] x [0, ny*dy]
    init_grid(0.0, in->dx * static_cast<Float>(NX), NX, 0.0, in->dy * static_cast<Float>(NY), NY, fs);
    fs.dx = in->dx; fs.dy = in->dy;
    cp_in(fs.visc, in->visc); cp_in(fs.p, in->p);
    cp_in(fs.p_jump_u_stag, in->p_jump_u); cp_in(fs.p_jump_v_stag, in->p_jump_v);
    cp_in(fs.old.rho_u_stag, in->rho_u_old); cp_in(fs.old.rho_v_stag, in->rho_v_old);
    cp_in(fs.old.U, in->U_old); cp_in(fs.old.V, in->V_old);
    cp_in(fs.curr.rho_u_stag, in->rho_u); cp_in(fs.curr.rho_v_stag, in->rho_v);
    cp_in(fs.curr.U, in->U); cp_in(fs.curr.V, in->V);
  }
  void store(fso_fs* out) const override {
    cp_out(fs.visc, out->visc); cp_out(fs.p, out->p);
    cp_out(fs.p_jump_u_stag, out->p_jump_u); cp_out(fs.p_jump_v_stag, out->p_jump_v);
    cp_out(fs.old.rho_u_stag, out->rho_u_old); cp_out(fs.old.rho_v_stag, out->rho_v_old);
    cp_out(fs.old.U, out->U_old); cp_out(fs.old.V, out->V_old);
    cp_out(fs.curr.rho_u_stag, out->rho_u); cp_out(fs.curr.rho_v_stag, out->rho_v);
    cp_out(fs.curr.U, out->U); cp_out(fs.curr.V, out->V);
  }
  // copies of the caller's profile tables (left/right: U at ym(j), V at y(j); bottom/top: U at x(i),
  // V at xm(i); all ghost inclusive) and the (side, component) slots they are published in
  std::vector<double> prof_[4][2];
  void publish_profiles() {
    for (int s = 0; s < 4; ++s)
      for (int q = 0; q < 2; ++q) {
        const bool lr = s < 2;
        const double h = lr ? fs.dy : fs.dx;
        // U on left/right and V on bottom/top live at cell centres of the side, the others at faces
        const bool centred = lr ? (q == 0) : (q == 1);
        g_prof[s][q] = ProfTable{prof_[s][q].empty() ? nullptr : prof_[s][q].data(),
                                 centred ? -0.5 * h : -h, h, static_cast<long>(prof_[s][q].size())};
      }
  }
  BCond_t<Float> make_bc(const fso_bconds* bc, int s) {
    const bool lr = s < 2;
    const size_t nu = static_cast<size_t>(lr ? NY + 2 : NX + 3), nv = static_cast<size_t>(lr ? NY + 3 : NX + 2);
    prof_[s][0].clear();
    prof_[s][1].clear();
    if (bc->kind[s] == FSO_BC_DIRICHLET) {
      Dirichlet<Float> d{.U = bc->u_val[s], .V = bc->v_val[s]};
      if (bc->u_prof[s]) { prof_[s][0].assign(bc->u_prof[s], bc->u_prof[s] + nu); d.U = PROF_FNS[s][0]; }
      if (bc->v_prof[s]) { prof_[s][1].assign(bc->v_prof[s], bc->v_prof[s] + nv); d.V = PROF_FNS[s][1]; }
      return d;
    }
    switch (bc->kind[s]) {
      case FSO_BC_NEUMANN:   return Neumann{.clipped = bc->clipped[s] != 0};
      case FSO_BC_PERIODIC:  return Periodic{};
      default:               return Symmetry{};
    }
  }
  void set_bconds(const fso_bconds* bc) override {
    bconds = FlowBConds<Float>{make_bc(bc, 0), make_bc(bc, 1), make_bc(bc, 2), make_bc(bc, 3)};
    publish_profiles();
  }
  double adjust_dt(double cfl, double dt_max) override { return ::adjust_dt(fs, cfl, dt_max); }
  void save_old_state() override { ::save_old_state(fs.curr, fs.old); }
  void calc_mid_time() override {
    ::calc_mid_time(fs.curr.U, fs.old.U);
    ::calc_mid_time(fs.curr.V, fs.old.V);
  }
  void calc_dmomdt(double* oU, double* oV) override {
    ::calc_dmomdt(fs, dU, dV);
    cp_out(dU, oU); cp_out(dV, oV);
  }
  void update_velocity(const double* iU, const double* iV, double dt) override {
    cp_in(dU, iU); cp_in(dV, iV);
    ::update_velocity(dU, dV, dt, fs);
  }
  void apply_bconds() override {
    publish_profiles();
    ::apply_velocity_bconds(fs, bconds, 0.0);
  }
  void calc_divergence(double* o) override {
    ::calc_divergence(fs.curr.U, fs.curr.V, fs.dx, fs.dy, div);
    cp_out(div, o);
  }
  void fill_operator(int psd, double* op) override {
    auto ls = std::make_unique<LinearSystem<Float, NX, NY, G, Layout::F>>();
    ls->fill_pressure_operator(fs, static_cast<PSDirichlet>(psd));
    std::memcpy(op, ls->op.get_data()->data(), sizeof(double) * 5 * static_cast<size_t>(ls->op.size()));
  }
  void fill_rhs(const double* idiv, double dt, int psd, double* rhs) override {
    auto ls = std::make_unique<LinearSystem<Float, NX, NY, G, Layout::F>>();
    cp_in(div, idiv);
    ls->fill_pressure_rhs(fs, div, dt, static_cast<PSDirichlet>(psd));
    std::memcpy(rhs, ls->rhs.get_data(), sizeof(double) * static_cast<size_t>(ls->rhs.size()));
  }
  void shift_pressure(double* io) override {
    cp_in(dp, io);
    ::shift_pressure_to_zero(fs.dx, fs.dy, dp);
    cp_out(dp, io);
  }
  void calc_drhodt(double* oU, double* oV) override {
    ::calc_drhodt(fs, dU, dV);
    cp_out(dU, oU); cp_out(dV, oV);
  }
  void rho_visc_vof(const double* vf) override {
    auto f = std::make_unique<Field2D<Float, NX, NY, G>>();
    cp_in(*f, vf);
    ::calc_rho(*f, fs);
    ::calc_visc(*f, fs);
  }
  void update_density(const double* iU, const double* iV, double dt) override {
    cp_in(dU, iU); cp_in(dV, iV);
    ::update_density(dU, dV, dt, fs);
  }
  // cell-field scratch for the two-phase inputs (vf, curv, interface_length)
  std::unique_ptr<Field2D<Float, NX, NY, G>> c_vf{}, c_curv{}, c_ilen{};
  std::unique_ptr<Field2D<Float, NX + 1, NY, G>> ib_u{};
  std::unique_ptr<Field2D<Float, NX, NY + 1, G>> ib_v{};
  void load_cells(const double* vf, const double* curv, const double* ilen) {
    if (!c_vf) {
      c_vf   = std::make_unique<Field2D<Float, NX, NY, G>>();
      c_curv = std::make_unique<Field2D<Float, NX, NY, G>>();
      c_ilen = std::make_unique<Field2D<Float, NX, NY, G>>();
    }
    cp_in(*c_vf, vf); cp_in(*c_curv, curv); cp_in(*c_ilen, ilen);
  }
  void pressure_jump(const double* vf, const double* curv, const double* ilen) override {
    load_cells(vf, curv, ilen);
    ::calc_pressure_jump(*c_vf, *c_curv, *c_ilen, fs);
  }
  void interpolate_uv(double* ui, double* vi) override {
    auto a = std::make_unique<Field2D<Float, NX, NY, G>>();
    ::interpolate_U(fs.curr.U, *a);
    cp_out(*a, ui);
    ::interpolate_V(fs.curr.V, *a);
    cp_out(*a, vi);
  }
  // to_npy(directory, fs), src/IO.hpp:231-269 (removes and re-creates the directory)
  int to_npy(const char* dir) override { return ::to_npy(std::string(dir), fs) ? 0 : 1; }
  // the drivers' DataWriter (src/VTKWriter.hpp with FS_DISABLE_HDF, src/IO.hpp:13-16) set up as in
  // test/PeriodicChannel.cpp:131-134: pressure, divergence, cell-centre velocity
  int write_vtk(const char* dir, double t) override {
    auto ui = std::make_unique<Field2D<Float, NX, NY, G>>();
    auto vi = std::make_unique<Field2D<Float, NX, NY, G>>();
    ::interpolate_U(fs.curr.U, *ui);
    ::interpolate_V(fs.curr.V, *vi);
    ::calc_divergence(fs.curr.U, fs.curr.V, fs.dx, fs.dy, div);
    DataWriter<Float, NX, NY, G> w(dir, &fs.x, &fs.y);
    w.add_scalar("pressure", &fs.p);
    w.add_scalar("divergence", &div);
    w.add_vector("velocity", ui.get(), vi.get());
    return w.write(t) ? 0 : 1;
  }
  void conserved(double* mass, double* mx, double* my) override {
    ::calc_conserved_quantities(fs, *mass, *mx, *my);
  }
  void ib_implicit_euler(const double* ibu, const double* ibv, double dt) override {
    auto a = std::make_unique<Field2D<Float, NX + 1, NY, G>>();
    auto b = std::make_unique<Field2D<Float, NX, NY + 1, G>>();
    cp_in(*a, ibu); cp_in(*b, ibv);
    ::correct_velocity_ib_implicit_euler(*a, *b, dt, fs);
  }
  void ib_semi_analytical(const double* iU, const double* iV, double dt, const double* ibu,
                          const double* ibv) override {
    auto a = std::make_unique<Field2D<Float, NX + 1, NY, G>>();
    auto b = std::make_unique<Field2D<Float, NX, NY + 1, G>>();
    cp_in(*a, ibu); cp_in(*b, ibv);
    cp_in(dU, iU); cp_in(dV, iV);
    ::update_velocity_ib_semi_analytical(dU, dV, dt, *a, *b, fs);
  }
  void solver_create(double tol, int max_iter, int psd) override {
    psd_ = psd;
    ps.reset();
    ps = std::make_unique<LinearSolver_StructHypre<Float, NX, NY, G>>(
        tol, max_iter, HypreSolver::PCG, HyprePrecond::PFMG, static_cast<PSDirichlet>(psd));
  }
  void solver_set_operator() override { ps->set_pressure_operator(fs); }
  int solver_solve(const double* idiv, double dt, double* odp, double* res) override {
    cp_in(div, idiv);
    Index it = 0;
    Float r  = 0;
    ps->set_pressure_rhs(fs, div, dt);
    ps->solve(dp, &r, &it);
    cp_out(dp, odp);
    if (res) *res = r;
    return static_cast<int>(it);
  }
  // The reference keeps these corrections in its drivers; the statements below are the drivers' own
  // (mode 1 test/LaminarChannel.cpp:55-66,201-205; mode 2 test/CustomDirichletChannel.cpp:61-72,197-202;
  // mode 3 test/PeriodicChannel.cpp:56-67,193-203), instantiated on this harness' fields.
  double flow_control(int mode, double total_flow, double* io) override {
    Float inflow = 0, outflow = 0;
    if (mode == 1) {
      for_each_a(fs.ym, [&](Index j) {
        inflow  += fs.curr.rho_u_stag(-G, j) * fs.curr.U(-G, j);
        outflow += fs.curr.rho_u_stag(NX + G, j) * fs.curr.U(NX + G, j);
      });
    } else {
      for_each_a(fs.ym, [&](Index j) {
        inflow  += fs.curr.rho_u_stag(-G, j) * fs.curr.U(-G, j) * fs.dy;
        outflow += fs.curr.rho_u_stag(NX + G, j) * fs.curr.U(NX + G, j) * fs.dy;
      });
    }
    if (io) { io[0] = inflow; io[1] = outflow; }
    const Float mass_error = outflow - inflow;
    if (mode == 3) {
      const Float inflow_error  = total_flow - inflow;
      const Float outflow_error = total_flow - outflow;
      for_each_a<Exec::Parallel>(fs.ym, [&](Index j) {
        fs.curr.U(-G, j)     += inflow_error / (fs.curr.rho_u_stag(-G, j) * fs.dy *
                                            static_cast<Float>(NY + 2 * G));
        fs.curr.U(NX + G, j) += outflow_error / (fs.curr.rho_u_stag(NX + G, j) * fs.dy *
                                                 static_cast<Float>(NY + 2 * G));
      });
    } else {
      for_each_a<Exec::Parallel>(fs.ym, [&](Index j) {
        fs.curr.U(NX + G, j) -=
            mass_error / (fs.curr.rho_u_stag(NX + G, j) * static_cast<Float>(NY + 2 * G));
      });
    }
    return mass_error;
  }
  // The sub-iteration loop of scaling/LidDrivenFlow.cpp:124-161 with the variable-density
  // correction of test/LaminarChannel.cpp:222-233 and its outflow fix (:201-205).
  // phase_secs (optional, 4 entries): momentum(mid-time+dmomdt+update+bc), div+rhs, solve, correct.
  int step(const fso_step_opts* o, double dt, double* res, double* phase) override {
    using clk = std::chrono::steady_clock;
    publish_profiles();
    auto secs = [](clk::time_point a, clk::time_point b) {
      return std::chrono::duration<double>(b - a).count();
    };
    ::save_old_state(fs.curr, fs.old);
    int iters = 0;
    Float r   = 0;
    for (int sub = 0; sub < o->num_subiter; ++sub) {
      auto t0 = clk::now();
      ::calc_mid_time(fs.curr.U, fs.old.U);
      ::calc_mid_time(fs.curr.V, fs.old.V);
      if (o->density_update) {  // the calls of test/StationaryDrop.cpp:274-293
        ::calc_drhodt(fs, dU, dV);
        ::update_density(dU, dV, dt, fs);
        ::apply_neumann_bconds(fs.curr.rho_u_stag);
        ::apply_neumann_bconds(fs.curr.rho_v_stag);
      }
      ::calc_dmomdt(fs, dU, dV);
      if (o->ib_mode) {  // the calls of scaling/LuchiniIBChannel.cpp:212-221
        if (!ib_u) {
          ib_u = std::make_unique<Field2D<Float, NX + 1, NY, G>>();
          ib_v = std::make_unique<Field2D<Float, NX, NY + 1, G>>();
        }
        cp_in(*ib_u, o->ib_u);
        cp_in(*ib_v, o->ib_v);
        if (o->ib_mode == 1) {
          ::update_velocity(dU, dV, dt, fs);
          ::correct_velocity_ib_implicit_euler(*ib_u, *ib_v, dt, fs);
        } else {
          ::update_velocity_ib_semi_analytical(dU, dV, dt, *ib_u, *ib_v, fs);
        }
      } else {
        for_each_i<Exec::Parallel>(fs.curr.U, [&](Index i, Index j) {
          fs.curr.U(i, j) =
              (fs.old.rho_u_stag(i, j) * fs.old.U(i, j) + dt * dU(i, j)) / fs.curr.rho_u_stag(i, j);
        });
        for_each_i<Exec::Parallel>(fs.curr.V, [&](Index i, Index j) {
          fs.curr.V(i, j) =
              (fs.old.rho_v_stag(i, j) * fs.old.V(i, j) + dt * dV(i, j)) / fs.curr.rho_v_stag(i, j);
        });
      }
      ::apply_velocity_bconds(fs, bconds, 0.0);
      if (o->outflow_fix) flow_control(o->outflow_fix, o->total_flow, nullptr);
      auto t1 = clk::now();
      ::calc_divergence(fs.curr.U, fs.curr.V, fs.dx, fs.dy, div);
      if (o->capillary) {  // the calls of test/StationaryDrop.cpp:331-353 (curvature is an input)
        load_cells(o->vf, o->curv, o->ilen);
        copy(fs.p_jump_u_stag, dU);
        copy(fs.p_jump_v_stag, dV);
        ::calc_pressure_jump(*c_vf, *c_curv, *c_ilen, fs);
        ::apply_neumann_bconds(fs.p_jump_u_stag);
        ::apply_neumann_bconds(fs.p_jump_v_stag);
        for_each_a<Exec::Parallel>(dU, [&](Index i, Index j) { dU(i, j) = fs.p_jump_u_stag(i, j) - dU(i, j); });
        for_each_a<Exec::Parallel>(dV, [&](Index i, Index j) { dV(i, j) = fs.p_jump_v_stag(i, j) - dV(i, j); });
        for_each_a<Exec::Parallel>(div, [&](Index i, Index j) {
          div(i, j) += dt * ((dU(i + 1, j) / fs.curr.rho_u_stag(i + 1, j) -
                              dU(i, j) / fs.curr.rho_u_stag(i, j)) /
                                 fs.dx +
                             (dV(i, j + 1) / fs.curr.rho_v_stag(i, j + 1) -
                              dV(i, j) / fs.curr.rho_v_stag(i, j)) /
                                 fs.dy);
        });
      }
      if (o->rebuild_operator_each_subiter) ps->set_pressure_operator(fs);
      ps->set_pressure_rhs(fs, div, dt);
      auto t2 = clk::now();
      Index it = 0;
      ps->solve(dp, &r, &it);
      iters += static_cast<int>(it);
      auto t3 = clk::now();
      ::shift_pressure_to_zero(fs.dx, fs.dy, dp);
      for_each_a<Exec::Parallel>(fs.p, [&](Index i, Index j) { fs.p(i, j) += dp(i, j); });
      for_each_i<Exec::Parallel>(fs.curr.U, [&](Index i, Index j) {
        fs.curr.U(i, j) -= (dp(i, j) - dp(i - 1, j)) / fs.dx * dt / fs.curr.rho_u_stag(i, j);
      });
      for_each_i<Exec::Parallel>(fs.curr.V, [&](Index i, Index j) {
        fs.curr.V(i, j) -= (dp(i, j) - dp(i, j - 1)) / fs.dy * dt / fs.curr.rho_v_stag(i, j);
      });
      auto t4 = clk::now();
      if (phase) {
        phase[0] += secs(t0, t1); phase[1] += secs(t1, t2);
        phase[2] += secs(t2, t3); phase[3] += secs(t3, t4);
      }
    }
    if (res) *res = r;
    return iters;
  }
  // mode 0: eval_flow_field_at(xm, ym, Ui, Vi, x, y) (src/Operators.hpp:118-161) with Ui, Vi from interpolate_U / _V;
  // mode 1: bilinear_interpolate(xm, ym, p, x, y) (:85-115) -> a
  void sample(int mode, int n, const double* px, const double* py, double* a, double* b) override {
    auto ui = std::make_unique<Field2D<Float, NX, NY, G>>();
    auto vi = std::make_unique<Field2D<Float, NX, NY, G>>();
    ::interpolate_U(fs.curr.U, *ui);
    ::interpolate_V(fs.curr.V, *vi);
    for (int k = 0; k < n; ++k) {
      if (mode == 0) {
        const auto uv = ::eval_flow_field_at(fs.xm, fs.ym, *ui, *vi, px[k], py[k]);
        a[k] = uv.first;
        b[k] = uv.second;
      } else {
        a[k] = ::bilinear_interpolate(fs.xm, fs.ym, fs.p, px[k], py[k]);
      }
    }
  }
  // the initial volume fraction of test/StationaryDrop.cpp:47-52,209-212: 16-point Gauss quadrature
  // (src/Quadrature.hpp:37-66) of the ellipse indicator over every cell incl. ghosts, then apply_neumann_bconds
  void vf_drop(double* out) override {
    constexpr Float CX = 0.5, CY = 0.5, R0 = 0.25;
    constexpr auto vof0 = [](Float x, Float y) {
      return static_cast<Float>(Igor::sqr(2.0 * (x - CX)) + Igor::sqr(y - CY) <= Igor::sqr(R0));
    };
    for_each_a<Exec::Parallel>(div, [&](Index i, Index j) {
      div(i, j) = quadrature(vof0, fs.x(i), fs.x(i + 1), fs.y(j), fs.y(j + 1)) / (fs.dx * fs.dy);
    });
    ::apply_neumann_bconds(div);
    cp_out(div, out);
  }
  // calc_ib_correction_shape (src/IB.hpp:45-107) for the two walls of scaling/LuchiniIBChannel.cpp:36-47,203-209:
  // solid below y0 and above y1
  void ib_channel_correction(double y0, double y1, double* ibu, double* ibv) override {
    const Float lx = fs.dx * static_cast<Float>(NX), ly = fs.dy * static_cast<Float>(NY);
    const Rect<Float> wall_bottom{.x = -1.0, .y = 0.0, .w = lx + 1.0, .h = y0};
    const Rect<Float> wall_top{.x = -1.0, .y = y1, .w = lx + 1.0, .h = ly - y1};
    auto a = std::make_unique<Field2D<Float, NX + 1, NY, G>>();
    auto b = std::make_unique<Field2D<Float, NX, NY + 1, G>>();
    fill(*a, 0.0);
    fill(*b, 0.0);
    ::calc_ib_correction_shape(wall_bottom, fs.dx, fs.dy, fs.x, fs.ym, *a);
    ::calc_ib_correction_shape(wall_bottom, fs.dx, fs.dy, fs.xm, fs.y, *b);
    ::calc_ib_correction_shape(wall_top, fs.dx, fs.dy, fs.x, fs.ym, *a);
    ::calc_ib_correction_shape(wall_top, fs.dx, fs.dy, fs.xm, fs.y, *b);
    cp_out(*a, ibu);
    cp_out(*b, ibv);
  }
  // integrate / L1_norm (src/Operators.hpp:51-76) and calc_grad_of_centered_points (:164-201) on a cell field
  double integrate_cell(const double* f, int include_ghost, int l1) override {
    cp_in(div, f);
    if (l1) return include_ghost ? ::L1_norm<true>(fs.dx, fs.dy, div) : ::L1_norm<false>(fs.dx, fs.dy, div);
    return include_ghost ? ::integrate<true>(fs.dx, fs.dy, div) : ::integrate<false>(fs.dx, fs.dy, div);
  }
  void grad_centered(const double* f, double* dfdx, double* dfdy) override {
    auto gx = std::make_unique<Field2D<Float, NX, NY, G>>();
    auto gy = std::make_unique<Field2D<Float, NX, NY, G>>();
    cp_in(div, f);
    ::calc_grad_of_centered_points(div, fs.dx, fs.dy, *gx, *gy);
    cp_out(*gx, dfdx);
    cp_out(*gy, dfdy);
  }
  void diagnostics(double* umax, double* vmax, double* divmax) override {
    ::calc_divergence(fs.curr.U, fs.curr.V, fs.dx, fs.dy, div);
    *umax = abs_max(fs.curr.U); *vmax = abs_max(fs.curr.V); *divmax = abs_max(div);
  }
};

SessionBase* make_session(int64_t nx, int64_t ny) {
#define X(A, B) \
  if (nx == (A) && ny == (B)) return new Session<(A), (B)>();
#include "ref_harness_sizes.inc"
#undef X
  return nullptr;
}

}  // namespace

extern "C" {

int fsref_supported(int64_t nx, int64_t ny) {
#define X(A, B) \
  if (nx == (A) && ny == (B)) return 1;
#include "ref_harness_sizes.inc"
#undef X
  return 0;
}
void fsref_set_threads(int n) { omp_set_num_threads(n); }
int fsref_max_threads(void) { return omp_get_max_threads(); }

void* fsref_create(int64_t nx, int64_t ny) { return make_session(nx, ny); }
void fsref_destroy(void* s) { delete static_cast<SessionBase*>(s); }
#define S static_cast<SessionBase*>(s)
void fsref_load(void* s, const fso_fs* in) { S->load(in); }
void fsref_store(void* s, fso_fs* out) { S->store(out); }
void fsref_set_bconds(void* s, const fso_bconds* bc) { S->set_bconds(bc); }
double fsref_adjust_dt(void* s, double cfl, double dt_max) { return S->adjust_dt(cfl, dt_max); }
void fsref_save_old_state(void* s) { S->save_old_state(); }
void fsref_calc_mid_time(void* s) { S->calc_mid_time(); }
void fsref_calc_dmomdt(void* s, double* dU, double* dV) { S->calc_dmomdt(dU, dV); }
void fsref_update_velocity(void* s, const double* dU, const double* dV, double dt) {
  S->update_velocity(dU, dV, dt);
}
void fsref_apply_velocity_bconds(void* s) { S->apply_bconds(); }
void fsref_calc_divergence(void* s, double* div) { S->calc_divergence(div); }
void fsref_fill_pressure_operator(void* s, int psd, double* op) { S->fill_operator(psd, op); }
void fsref_fill_pressure_rhs(void* s, const double* div, double dt, int psd, double* rhs) {
  S->fill_rhs(div, dt, psd, rhs);
}
void fsref_shift_pressure_to_zero(void* s, double* dp) { S->shift_pressure(dp); }
void fsref_calc_drhodt(void* s, double* dru, double* drv) { S->calc_drhodt(dru, drv); }
void fsref_calc_rho_visc_vof(void* s, const double* vf) { S->rho_visc_vof(vf); }
void fsref_solver_create(void* s, double tol, int max_iter, int psd) {
  S->solver_create(tol, max_iter, psd);
}
void fsref_solver_set_operator(void* s) { S->solver_set_operator(); }
int fsref_solver_solve(void* s, const double* div, double dt, double* dp, double* res) {
  return S->solver_solve(div, dt, dp, res);
}
int fsref_step(void* s, const fso_step_opts* o, double dt, double* res, double* phase_secs) {
  return S->step(o, dt, res, phase_secs);
}
void fsref_diagnostics(void* s, double* umax, double* vmax, double* divmax) {
  S->diagnostics(umax, vmax, divmax);
}
void fsref_update_density(void* s, const double* dru, const double* drv, double dt) {
  S->update_density(dru, drv, dt);
}
void fsref_calc_pressure_jump(void* s, const double* vf, const double* curv, const double* ilen) {
  S->pressure_jump(vf, curv, ilen);
}
void fsref_interpolate_uv(void* s, double* ui, double* vi) { S->interpolate_uv(ui, vi); }
void fsref_conserved(void* s, double* mass, double* mx, double* my) { S->conserved(mass, mx, my); }
int fsref_to_npy(void* s, const char* dir) { return S->to_npy(dir); }
void fsref_ib_channel_correction(void* s, double y0, double y1, double* ibu, double* ibv) {
  S->ib_channel_correction(y0, y1, ibu, ibv);
}
void fsref_vf_drop(void* s, double* vf) { S->vf_drop(vf); }
void fsref_sample(void* s, int mode, int n, const double* px, const double* py, double* a, double* b) {
  S->sample(mode, n, px, py, a, b);
}
// Monitor (src/Monitor.hpp) driven with nrows x (nf Float + ni Index) values: the log file as the drivers write it
int fsref_monitor_log(const char* filename, int nf, const char* const* fnames, int ni, const char* const* inames,
                      int nrows, const double* fvals, const int64_t* ivals) {
  std::vector<Float> fv(static_cast<size_t>(nf));
  std::vector<Index> iv(static_cast<size_t>(ni));
  Monitor<Float> monitor(filename);
  // the drivers interleave Float and Index columns; here: all Float columns but the last, the Index columns, the last Float
  for (int k = 0; k + 1 < nf; ++k) monitor.add_variable(&fv[static_cast<size_t>(k)], fnames[k]);
  for (int k = 0; k < ni; ++k) monitor.add_variable(&iv[static_cast<size_t>(k)], inames[k]);
  if (nf > 0) monitor.add_variable(&fv[static_cast<size_t>(nf - 1)], fnames[nf - 1]);
  for (int r = 0; r < nrows; ++r) {
    for (int k = 0; k < nf; ++k) fv[static_cast<size_t>(k)] = fvals[r * nf + k];
    for (int k = 0; k < ni; ++k) iv[static_cast<size_t>(k)] = ivals[r * ni + k];
    monitor.write();
  }
  return 0;
}
// should_save(t, dt, dt_write, t_end), src/IO.hpp:97-108 (keeps a function-static last save time)
int fsref_should_save(double t, double dt, double dt_write, double t_end) {
  return ::should_save(t, dt, dt_write, t_end) ? 1 : 0;
}
double fsref_integrate_cell(void* s, const double* f, int include_ghost, int l1) {
  return S->integrate_cell(f, include_ghost, l1);
}
void fsref_grad_centered(void* s, const double* f, double* dfdx, double* dfdy) { S->grad_centered(f, dfdx, dfdy); }
int fsref_write_vtk(void* s, const char* dir, double t) { return S->write_vtk(dir, t); }
double fsref_flow_control(void* s, int mode, double total_flow, double* io) {
  return S->flow_control(mode, total_flow, io);
}
void fsref_ib_implicit_euler(void* s, const double* ibu, const double* ibv, double dt) {
  S->ib_implicit_euler(ibu, ibv, dt);
}
void fsref_ib_semi_analytical(void* s, const double* dU, const double* dV, double dt,
                              const double* ibu, const double* ibv) {
  S->ib_semi_analytical(dU, dV, dt, ibu, ibv);
}
#undef S
}
