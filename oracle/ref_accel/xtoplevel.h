// TEST INFRASTRUCTURE — stands in for the un-vendored Xilinx HLS driver header the reference's AcceleratorHandler.h
// includes (src/backend/AcceleratorHandler.h:8; generated into ${HLS_PROJECT}/impl/ip/drivers/toplevel_v1_0/src by
// the FPGA tool flow, src/backend/CMakeLists.txt:25-30).  The reference's class declaration is compiled UNMODIFIED
// against this; the one member of that type (xToplevelInstance) carries what the B200-backed implementation in
// AcceleratorHandler_slamgpu.cpp needs between setN() and start().
#pragma once
#include <stdint.h>
#include <sys/types.h>

typedef struct {
    uint32_t n;      // landmarks in the record (setN)
    uint32_t done;   // the last start() completed
    uint32_t calls;
} XToplevel;
