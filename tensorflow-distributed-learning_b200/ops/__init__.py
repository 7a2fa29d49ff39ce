"""ctypes bindings of the sm_100a kernels (libtfdl_kernels.so)."""
