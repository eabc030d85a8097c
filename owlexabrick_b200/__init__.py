"""owlexabrick_b200 — Blackwell-native ExaBricks rendering hot path.

`capi` binds libexabrick_b200.so (built in-tree by `python -m owlexabrick_b200.build`);
`renderer` mirrors the reference's host classes on top of it; `synth` makes synthetic AMR
inputs.  Importing `capi`/`renderer` raises ImportError when the library is missing.
"""
__all__ = ["capi", "renderer", "synth", "build"]
