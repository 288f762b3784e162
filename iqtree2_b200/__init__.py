"""B200-native tree-likelihood engine behind IQ-TREE 2's kernel-dispatch API.

csrc/      hand-written CUDA for sm_100a + the C ABI of include/b200lk.h
capi.py    ctypes binding of that ABI
workload.py synthetic inputs (trees, eigen-systems, alignments)
"""
