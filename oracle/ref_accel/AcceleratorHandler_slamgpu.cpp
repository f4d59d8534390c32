// TEST INFRASTRUCTURE — the literal accelerator seam, compiled.
//
// The reference selects its FPGA Jacobian core at compile time with -DJACOBIAN_ACCELERATOR=1: computeJacobians
// (core.cpp:586-664) then marshals one particle into the buffer AcceleratorHandler::getMemoryPointer() returns, calls
// setN / start, spins on isDone and reads the 16 floats per landmark back.  This file implements the reference's OWN
// class (src/backend/AcceleratorHandler.h:10-33, compiled unmodified) on top of the B200 library instead of /dev/mem
// and the HLS driver (AcceleratorHandler.cpp:16-66): start() hands the record to slamgpu_compute_jacobians, whose
// in / out layout is that record.  oracle/Makefile links it with the reference's unmodified core, algorithms, wrappers
// and application into oracle/_ref/slam-backend-ref-accel; tests/test_gpu_accel_seam.py runs the reference's own
// FastSLAM2 through it and holds the trace against the CPU reference's.
#include <cstdio>
#include <cstdlib>
#include <stdexcept>

#include "src/backend/AcceleratorHandler.h"
#include "include/slamgpu.h"

static const size_t kRecordBytes = 256 * 1024;  // the OCM window of the FPGA design: n <= 2978

AcceleratorHandler::AcceleratorHandler() {
    ocm = calloc(1, kRecordBytes);
    if (!ocm) throw std::runtime_error("AcceleratorHandler: no memory for the record buffer");
    xToplevelInstance.n = 0;
    xToplevelInstance.done = 1;
    xToplevelInstance.calls = 0;
}

AcceleratorHandler::~AcceleratorHandler() {
    if (getenv("SLAMREF_ACCEL_STATS")) fprintf(stderr, "AcceleratorHandler: %u records through slamgpu\n", xToplevelInstance.calls);
    free(ocm);
}

void *AcceleratorHandler::getMemoryPointer() { return ocm; }

void AcceleratorHandler::setN(uint32_t n) { xToplevelInstance.n = n; }

void AcceleratorHandler::start() {
    const uint32_t n = xToplevelInstance.n;
    xToplevelInstance.done = 0;
    if ((7 + 6 * (size_t) n + 16 * (size_t) n) * sizeof(float) > kRecordBytes)
        throw std::runtime_error("AcceleratorHandler: record exceeds the 256 KB window");
    float *rec = (float *) ocm;
    // in: xv[3], R(0..3), n x {xf[2], Pf(0..3)}; out, right behind: n x {zp[2], Hf[4], Hv[6], Sf[4]}
    const int rc = slamgpu_compute_jacobians(nullptr, rec, (int32_t) n, rec + 7 + 6 * (size_t) n);
    if (rc != SLAMGPU_OK) throw std::runtime_error(std::string("slamgpu_compute_jacobians: ") + slamgpu_last_error());
    xToplevelInstance.calls++;
    xToplevelInstance.done = 1;
}

uint AcceleratorHandler::isDone() { return xToplevelInstance.done; }
