// Force-included when compiling the reference's .cu files for sm_100a: legacy warp shuffles were removed
// for sm_70+, so map them to the _sync forms with a full mask (the reference kernels are warp-synchronous
// and launch full warps).  No reference source is modified or copied.
#pragma once
#ifndef __shfl_xor
#define __shfl_xor(v, m) __shfl_xor_sync(0xffffffffu, (v), (m))
#endif
