"""Host-side O(K^2)/O(D^2) helpers of the engine (parameter draws, small-matrix algebra).  All O(N^2) work is in libgk."""
