"""B200-native hot path for mrankitvish/ai-agent-kubectl.

``strings``  -- validator / parser / sanitiser / cache-key hash on the GPU (k9)
``_lib``     -- ctypes binding of libkx_b200.so (include/kx.h)
``synth``    -- synthetic workloads named by BASELINE.json

The directory name contains a hyphen; import it as ``kubectl_b200`` (the alias
module at the repo root)."""
__all__ = ["strings", "synth", "_lib", "build"]
