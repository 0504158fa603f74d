"""bramble_b200: B200-native projection engine (genome -> transcriptome coordinates) behind a C ABI.

Host-side Python mirrors of the reference's ingest (pack.py), synthetic inputs (synth.py) and a thin
ctypes wrapper of the CUDA engine (engine.py / capi.py).  No CPU implementation of the hot path lives here."""
__all__ = ["capi", "pack", "synth", "engine"]
