// Host-emulation stub of the NVTX header: ranges are no-ops.
#pragma once
inline int nvtxRangePushA(const char*) { return 0; }
inline int nvtxRangePop() { return 0; }
