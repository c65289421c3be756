"""B200-native grasp-hypothesis hot path of gwding/agile_grasp2.

csrc/      hand-written CUDA (sm_100a) + the C ABI of include/ag2.h  -> csrc/libag2.so
capi.py    ctypes binding of that ABI (plumbing for tests, bench and the multi-GPU host)
sharded.py one-process-per-GPU sample sharding over torch.distributed
scenes.py  seeded synthetic input clouds (bench / tests only)
"""
__all__ = ["capi", "scenes", "sharded"]
