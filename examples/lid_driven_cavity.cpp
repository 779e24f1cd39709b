// examples/lid_driven_cavity.cpp -- the reference's scaling/LidDrivenFlow.cpp time loop
// (:100-174, constants :16-42) written against include/shim/FluidSolverGPU.hpp: same calls, same
// order; every operation runs on the B200 through libfsgpu.so.  Prints one line per step and a
// checksum the GPU test compares with the Python binding's run of the same case.
//
//   g++ -std=c++20 -O2 -Iinclude examples/lid_driven_cavity.cpp -Lfluidsolver_b200 -lfsgpu
//       -Wl,-rpath,$PWD/fluidsolver_b200 -o examples/lid_driven_cavity   (one line)
//   ./examples/lid_driven_cavity [steps]
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>

#include "shim/FluidSolverGPU.hpp"

using Float = double;
#ifndef CAVITY_N
#define CAVITY_N 256
#endif
constexpr Index NX = CAVITY_N, NY = CAVITY_N, NGHOST = 1;
constexpr Float X_MIN = 0.0, X_MAX = 1.0, Y_MIN = 0.0, Y_MAX = 1.0;
constexpr Float DT_MAX = 1e-1, CFL_MAX = 0.9, U_TOP = 1.0, VISC = 1e-3, RHO = 1.0;
constexpr int PRESSURE_MAX_ITER = 50;
constexpr Float PRESSURE_TOL    = 1e-6;
constexpr Index NUM_SUBITER     = 5;

int main(int argc, char** argv) {
  const int nsteps = argc > 1 ? std::atoi(argv[1]) : 10;
  const FlowBConds<Float> bconds{
      .left   = Dirichlet<Float>{.U = 0.0, .V = 0.0},
      .right  = Dirichlet<Float>{.U = 0.0, .V = 0.0},
      .bottom = Dirichlet<Float>{.U = 0.0, .V = 0.0},
      .top    = Dirichlet<Float>{.U = U_TOP, .V = 0.0},
  };
  FS<Float, NX, NY, NGHOST> fs{.visc_gas = VISC, .visc_liquid = VISC, .rho_gas = RHO, .rho_liquid = RHO};
  init_grid(X_MIN, X_MAX, NX, Y_MIN, Y_MAX, NY, fs);

  Field2D<Float, NX, NY, NGHOST> div{};
  Field2D<Float, NX + 1, NY, NGHOST> drhoUdt{};
  Field2D<Float, NX, NY + 1, NGHOST> drhoVdt{};
  Field2D<Float, NX, NY, NGHOST> delta_p{};

  calc_rho(fs);
  calc_visc(fs);
  LinearSolver_StructHypre<Float, NX, NY, NGHOST> ps(PRESSURE_TOL, PRESSURE_MAX_ITER);
  ps.set_pressure_operator(fs);

  // zero initial flow: the host mirrors are zero-initialised, send them once
  fs.to_device();
  Float t = 0.0;
  apply_velocity_bconds(fs, bconds, t);

  for (int step = 0; step < nsteps; ++step) {
    const Float dt = adjust_dt(fs, CFL_MAX, DT_MAX);
    save_old_state(fs.curr, fs.old);
    Index pressure_iter = 0;
    Float pressure_res  = 0.0;
    for (Index sub_iter = 0; sub_iter < NUM_SUBITER; ++sub_iter) {
      calc_mid_time(fs.curr.U, fs.old.U);
      calc_mid_time(fs.curr.V, fs.old.V);
      calc_dmomdt(fs, drhoUdt, drhoVdt);
      update_velocity(drhoUdt, drhoVdt, dt, fs);
      apply_velocity_bconds(fs, bconds, t);
      calc_divergence(fs.curr.U, fs.curr.V, fs.dx, fs.dy, div);
      Index local_pressure_iter = 0;
      ps.set_pressure_rhs(fs, div, dt);
      ps.solve(delta_p, &pressure_res, &local_pressure_iter);
      pressure_iter += local_pressure_iter;
      shift_pressure_to_zero(fs.dx, fs.dy, delta_p);
      correct_velocity(fs, delta_p, dt);
    }
    t += dt;
    calc_divergence(fs.curr.U, fs.curr.V, fs.dx, fs.dy, div);
    std::printf("step %d t %.6e dt %.6e U_max %.12e V_max %.12e div_max %.3e p_iter %lld p_res %.3e\n", step, t,
                dt, abs_max(fs.curr.U), abs_max(fs.curr.V), abs_max(div), (long long)pressure_iter, pressure_res);
  }
  fs.to_host();
  double su = 0.0, sv = 0.0, sp = 0.0;
  for (Index i = 0; i <= NX; ++i)
    for (Index j = 0; j < NY; ++j) su += fs.curr.U(i, j) * fs.curr.U(i, j);
  for (Index i = 0; i < NX; ++i)
    for (Index j = 0; j <= NY; ++j) sv += fs.curr.V(i, j) * fs.curr.V(i, j);
  for (Index i = 0; i < NX; ++i)
    for (Index j = 0; j < NY; ++j) sp += (fs.p(i, j) - fs.p(0, 0)) * (fs.p(i, j) - fs.p(0, 0));
  std::printf("checksum N %lld steps %d U2 %.15e V2 %.15e P2 %.15e\n", (long long)NX, nsteps, su, sv, sp);
  return 0;
}
