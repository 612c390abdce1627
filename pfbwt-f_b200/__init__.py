"""pfbwt-f_b200: B200-native prefix-free-parsing BWT builder (drop-in for alshai/pfbwt-f's hot path).
The product is libpfbwt_b200.so (CUDA, sm_100a) + the C++ drivers in bin/; this package only holds the
ctypes view of the C ABI used by the tests and the benchmark."""
