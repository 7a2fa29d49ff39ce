"""Host runtime (libtfdl_runtime.so): TCP control plane, ring all-reduce, data helpers, checkpoint IO."""
