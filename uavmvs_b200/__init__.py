"""uavmvs_b200 -- B200-native visibility-and-reconstructability evaluation
(the libs/eval + libs/cacc hot path of nmoehrle/uavmvs) behind a C ABI.

  uavmvs_b200.api     host-side mirror of the reference's cacc::/eval interface
                      (ctypes over libuavmvs_b200.so; needs a CUDA device)
  uavmvs_b200.scenes  seeded synthetic inputs (numpy, host only)
"""
__all__ = ["api", "scenes"]
