"""WaveNet vocoder: host-side mirror of the reference's ``wavenet_vocoder`` package for the
incremental mel->audio synthesis path, over the ``libwavenet_b200.so`` C ABI (sm_100a CUDA)."""
