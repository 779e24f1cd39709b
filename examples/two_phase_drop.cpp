// examples/two_phase_drop.cpp -- the sub-iteration of the reference's two-phase drivers
// (test/StationaryDrop.cpp:253-384, constants :28-72) written against include/shim/FluidSolverGPU.hpp
// with a FROZEN volume-fraction field: VOF advection, interface reconstruction and the curvature
// estimators are outside this library's scope, so vf, curv and
// interface_length are host-filled inputs here (a lid drives the flow so the solve has work to do).
// Prints a checksum the GPU test compares with the Python binding's run of the same case.
//
//   ./examples/two_phase_drop [steps]
#include <cmath>
#include <cstdio>
#include <cstdlib>

#include "shim/FluidSolverGPU.hpp"

using Float = double;
constexpr Index NX = 128, NY = 128, NGHOST = 1;
constexpr Float DT_MAX = 1e-1, CFL_MAX = 0.9, VISC = 1e-3, RHO_G = 1.0, RHO_L = 1000.0, SIGMA = 0.05;
constexpr Index NUM_SUBITER = 3;

int main(int argc, char** argv) {
  const int nsteps = argc > 1 ? std::atoi(argv[1]) : 2;
  const FlowBConds<Float> bconds{
      .left   = Dirichlet<Float>{.U = 0.0, .V = 0.0},
      .right  = Dirichlet<Float>{.U = 0.0, .V = 0.0},
      .bottom = Dirichlet<Float>{.U = 0.0, .V = 0.0},
      .top    = Dirichlet<Float>{.U = 1.0, .V = 0.0},
  };
  FS<Float, NX, NY, NGHOST> fs{.visc_gas = VISC, .visc_liquid = VISC, .rho_gas = RHO_G, .rho_liquid = RHO_L};
  fs.sigma = SIGMA;
  init_grid(0.0, 1.0, NX, 0.0, 1.0, NY, fs);

  Field2D<Float, NX, NY, NGHOST> vf{}, curv{}, interface_length{}, div{}, delta_p{}, Ui{}, Vi{}, rhoi{};
  Field2D<Float, NX + 1, NY, NGHOST> drho_u_stagdt{}, drhoUdt{};
  Field2D<Float, NX, NY + 1, NGHOST> drho_v_stagdt{}, drhoVdt{};

  // smoothed indicator of the ellipse (2(x-1/2))^2 + (y-1/2)^2 <= 0.25^2 (test/StationaryDrop.cpp:46-52)
  for (Index i = -1; i <= NX; ++i)
    for (Index j = -1; j <= NY; ++j) {
      const Float x = fs.xm(i), y = fs.ym(j);
      const Float d = std::sqrt(4.0 * (x - 0.5) * (x - 0.5) + (y - 0.5) * (y - 0.5)) - 0.25;
      Float f       = 0.5 - d / (1.5 * fs.dx);
      f             = f < 0.0 ? 0.0 : (f > 1.0 ? 1.0 : f);
      vf(i, j)      = f;
      curv(i, j)    = 4.0;
      interface_length(i, j) = (f > 1e-8 && f < 1.0 - 1e-8) ? fs.dx : 0.0;
    }
  fs.to_device();
  calc_rho(vf, fs);
  calc_visc(vf, fs);
  save_old_density(fs.curr, fs.old);
  LinearSolver_StructHypre<Float, NX, NY, NGHOST> ps(1e-10, 200);
  apply_velocity_bconds(fs, bconds, 0.0);

  for (int step = 0; step < nsteps; ++step) {
    const Float dt = adjust_dt(fs, CFL_MAX, DT_MAX);
    save_old_state(fs.curr, fs.old);
    Index p_iter = 0;
    Float p_res  = 0.0;
    for (Index sub_iter = 0; sub_iter < NUM_SUBITER; ++sub_iter) {
      calc_mid_time(fs.curr.U, fs.old.U);
      calc_mid_time(fs.curr.V, fs.old.V);
      calc_drhodt(fs, drho_u_stagdt, drho_v_stagdt);
      update_density(drho_u_stagdt, drho_v_stagdt, dt, fs);
      apply_neumann_bconds(fs.curr.rho_u_stag);
      apply_neumann_bconds(fs.curr.rho_v_stag);
      calc_dmomdt(fs, drhoUdt, drhoVdt);
      update_velocity(drhoUdt, drhoVdt, dt, fs);
      apply_velocity_bconds(fs, bconds, 0.0);
      calc_divergence(fs.curr.U, fs.curr.V, fs.dx, fs.dy, div);
      if (step == 0 && sub_iter == 0) calc_pressure_jump(vf, curv, interface_length, fs);  // uploads the inputs
      add_capillary_divergence(fs, div, dt);
      Index local_p_iter = 0;
      ps.set_pressure_operator(fs);
      ps.set_pressure_rhs(fs, div, dt);
      ps.solve(delta_p, &p_res, &local_p_iter);
      p_iter += local_p_iter;
      shift_pressure_to_zero(fs.dx, fs.dy, delta_p);
      correct_velocity(fs, delta_p, dt);
    }
    interpolate_U(fs.curr.U, Ui);
    interpolate_V(fs.curr.V, Vi);
    interpolate_UV_staggered_field(fs.curr.rho_u_stag, fs.curr.rho_v_stag, rhoi);
    calc_divergence(fs.curr.U, fs.curr.V, fs.dx, fs.dy, div);
    Float mass = 0, mom_x = 0, mom_y = 0;
    calc_conserved_quantities(fs, mass, mom_x, mom_y);
    std::printf("step %d dt %.6e U_max %.12e div_max %.3e p_iter %lld p_res %.3e mass %.15e mom_x %.15e\n", step, dt,
                abs_max(fs.curr.U), abs_max(div), (long long)p_iter, p_res, mass, mom_x);
  }
  fs.to_host();
  rhoi.download();
  double su = 0.0, sv = 0.0, sr = 0.0;
  for (Index i = 0; i <= NX; ++i)
    for (Index j = 0; j < NY; ++j) su += fs.curr.U(i, j) * fs.curr.U(i, j);
  for (Index i = 0; i < NX; ++i)
    for (Index j = 0; j <= NY; ++j) sv += fs.curr.V(i, j) * fs.curr.V(i, j);
  for (Index i = 0; i < NX; ++i)
    for (Index j = 0; j < NY; ++j) sr += rhoi(i, j);
  std::printf("checksum N %lld steps %d U2 %.15e V2 %.15e RHOI %.15e\n", (long long)NX, nsteps, su, sv, sr);
  return 0;
}
