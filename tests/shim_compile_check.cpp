// tests/shim_compile_check.cpp -- compile-and-link check of include/shim/FluidSolverGPU.hpp (CPU suite):
// every call of the reference's surface that the shim re-creates is instantiated once, in the shape the
// reference drivers use it (scaling/LidDrivenFlow.cpp, test/LaminarChannel.cpp, test/StationaryDrop.cpp,
// test/Couette.cpp).  Nothing runs without a GPU: main() returns before the first device call unless
// an argument is given.
#include "shim/FluidSolverGPU.hpp"

using Float = double;
constexpr Index NX = 32, NY = 16, NGHOST = 1;

static Float inlet(Float y, Float /*t*/) { return 6.0 * y * (1.0 - y); }

int main(int argc, char**) {
  if (argc < 2) return 0;
  const FlowBConds<Float> bconds{
      .left   = Dirichlet<Float>{.U = &inlet, .V = 0.0},
      .right  = Neumann{.clipped = true},
      .bottom = Symmetry{},
      .top    = Dirichlet<Float>{.U = 0.0, .V = 0.0},
  };
  const FlowBConds<Float> periodic{Periodic{}, Periodic{}, Periodic{}, Periodic{}};
  FS<Float, NX, NY, NGHOST> fs{.visc_gas = 1e-3, .visc_liquid = 1e-3, .rho_gas = 1.0, .rho_liquid = 1000.0};
  fs.sigma = 0.05;
  init_grid(0.0, 2.0, NX, 0.0, 1.0, NY, fs);
  Field2D<Float, NX, NY, NGHOST> vf{}, curv{}, ilen{}, div{}, delta_p{}, Ui{}, Vi{}, rhoi{};
  Field2D<Float, NX + 1, NY, NGHOST> drhoUdt{}, drho_u{}, ib_u{};
  Field2D<Float, NX, NY + 1, NGHOST> drhoVdt{}, drho_v{}, ib_v{};
  for_each_a<Exec::Parallel>(vf, [&](Index i, Index j) { vf(i, j) = fs.xm(i) < 1.0 ? 1.0 : 0.0; });
  for_each_i<Exec::Parallel>(fs.curr.U, [&](Index i, Index j) { fs.curr.U(i, j) = inlet(fs.ym(j), 0.0); });
  for_each_a(fs.ym, [&](Index j) { (void)fs.ym(j); });
  for_each<-1, NX + 1, 0, NY, Exec::Parallel>([&](Index i, Index j) { (void)fs.visc(i, j); });
  fs.to_device();
  calc_rho(fs);
  calc_visc(fs);
  calc_rho(vf, fs);
  calc_visc(vf, fs);
  apply_neumann_bconds(fs.curr.rho_u_stag);
  save_old_density(fs.curr, fs.old);
  save_old_velocity(fs.curr, fs.old);
  LinearSolver_StructHypre<Float, NX, NY, NGHOST> ps(1e-6, 50, HypreSolver::PCG, HyprePrecond::PFMG, PSDirichlet::RIGHT);
  ps.set_pressure_operator(fs);
  apply_velocity_bconds(fs, bconds, 0.0);
  apply_velocity_bconds(fs, periodic);
  const Float dt = adjust_dt(fs, 0.9, 0.1);
  save_old_state(fs.curr, fs.old);
  calc_mid_time(fs.curr.U, fs.old.U);
  calc_mid_time(fs.curr.V, fs.old.V);
  calc_drhodt(fs, drho_u, drho_v);
  update_density(drho_u, drho_v, dt, fs);
  calc_dmomdt(fs, drhoUdt, drhoVdt);
  update_velocity(drhoUdt, drhoVdt, dt, fs);
  update_velocity_ib_semi_analytical(drhoUdt, drhoVdt, dt, ib_u, ib_v, fs);
  correct_velocity_ib_implicit_euler(ib_u, ib_v, dt, fs);
  calc_divergence(fs.curr.U, fs.curr.V, fs.dx, fs.dy, div);
  calc_pressure_jump(vf, curv, ilen, fs);
  add_capillary_divergence(fs, div, dt);
  Float res = 0.0;
  Index it  = 0;
  ps.set_pressure_rhs(fs, div, dt);
  ps.solve(delta_p, &res, &it);
  shift_pressure_to_zero(fs.dx, fs.dy, delta_p);
  correct_velocity(fs, delta_p, dt);
  interpolate_U(fs.curr.U, Ui);
  interpolate_V(fs.curr.V, Vi);
  interpolate_UV_staggered_field(fs.curr.rho_u_stag, fs.curr.rho_v_stag, rhoi);
  Float mass = 0, mx = 0, my = 0;
  calc_conserved_quantities(fs, mass, mx, my);
  const Float m = abs_max(fs.curr.U) + abs_max<true>(div) + mass + mx + my;
  Field2D<Float, NX, NY, NGHOST> dvfdx{}, dvfdy{};
  calc_grad_of_centered_points(vf, fs.dx, fs.dy, dvfdx, dvfdy);
  const Float vol = integrate<true>(fs.dx, fs.dy, vf) + L1_norm(fs.dx, fs.dy, div);
  const auto [us, vs] = eval_flow_field_at(fs.xm, fs.ym, Ui, Vi, 0.7, 0.3);
  if (!(vol + us + vs + bilinear_interpolate(fs.xm, fs.ym, fs.p, 0.7, 0.3) == vol + us + vs + bilinear_interpolate(fs.xm, fs.ym, fs.p, 0.7, 0.3))) return 3;
  DataWriter<Float, NX, NY, NGHOST> data_writer("/tmp/fsgpu_shim_check", &fs.x, &fs.y);
  data_writer.add_scalar("pressure", &fs.p);
  data_writer.add_scalar("divergence", &div);
  data_writer.add_vector("velocity", &Ui, &Vi);
  const bool wrote = to_npy("/tmp/fsgpu_shim_check", fs) && data_writer.write(0.0) &&
                     to_npy("/tmp/fsgpu_shim_check/ym.npy", fs.ym) && to_npy("/tmp/fsgpu_shim_check/div.npy", div);
  if (!wrote || has_nan_or_inf(delta_p) || has_nan(fs.curr.U) || has_inf(div) || !div.is_valid_index(-1, NY)) return 2;
  fill(fs.p, 0.0);
  fs.to_host();
  return m == m ? 0 : 1;
}
