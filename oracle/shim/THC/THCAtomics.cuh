// TEST INFRASTRUCTURE ONLY. Empty stand-in: the reference's TorchSDF kernel includes <THC/THCAtomics.cuh>
// (unbatched_triangle_distance_cuda.cu:20) but uses nothing from it, and torch >= 2 no longer ships the header.
#pragma once
