"""lpcnet_b200: B200-native LPCNet synthesis engine (CUDA kernels behind the reference's C API).

The product is liblpcnet_b200.so (lpcnet_b200/csrc, C-ABI in include/lpcnet_b200.h); this
package holds the build recipe, the ctypes host mirror and the weight-blob tools."""
