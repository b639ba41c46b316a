// b2s_api.h -- internal glue between the C ABI (b2s_api.cu) and the stage kernels.
#pragma once
#include <cuda_runtime.h>
#include "../../include/b200sim.h"
#include "b2s_types.h"

void b2s_count_launch();

// stage launchers (each is one kernel launch on `st`)
void launch_position(const ModelDev& m, const DataDev& d, cudaStream_t st);
void launch_static_frames(const ModelDev& m, const DataDev& d, cudaStream_t st);
void launch_dynamics(const ModelDev& m, const DataDev& d, cudaStream_t st);
void launch_collision(const ModelDev& m, const DataDev& d, cudaStream_t st);
void launch_constraint(const ModelDev& m, const DataDev& d, cudaStream_t st);
void launch_efc_jacobian(const ModelDev& m, const DataDev& d, float* J, cudaStream_t st);
void launch_solve(const ModelDev& m, const DataDev& d, cudaStream_t st, cudaStream_t side = nullptr);
bool solve_has_second_launch(const ModelDev& m, const DataDev& d);
void launch_sensors(const ModelDev& m, const DataDev& d, cudaStream_t st);
void launch_integrate(const ModelDev& m, const DataDev& d, cudaStream_t st);

// dynamic shared memory each stage needs for this model (bytes per CTA) and its entry point,
// so the host can raise cudaFuncAttributeMaxDynamicSharedMemorySize once per model.
size_t smem_position(const ModelDev& m);  const void* fn_position();
size_t smem_dynamics(const ModelDev& m);  const void* fn_dynamics(const ModelDev& m);
size_t smem_collision(const ModelDev& m); const void* fn_collision(const ModelDev& m);
size_t smem_constraint(const ModelDev& m); const void* fn_constraint();
size_t smem_solve(const ModelDev& m, const DataDev& d); const void* fn_solve(const ModelDev& m);
int solve_prepare_all(const ModelDev& m, const DataDev& d, size_t* need, int device);
int b2s_raise_smem(int device, const void* fn, size_t smem);   // raise (never lower) a kernel's dynamic shared-memory limit
size_t smem_sensors(const ModelDev& m, const DataDev& d);   const void* fn_sensors();
size_t smem_integrate(const ModelDev& m); const void* fn_integrate(const ModelDev& m);
