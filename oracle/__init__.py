"""ORACLE = test infrastructure.  CPU restatements of the reference's algorithms for the decode hot path, plus loaders
for the reference's own code where it can run (oracle/_ref: the reference CPU ops compiled unmodified; refload: the
reference's Python pack/quant helpers imported by file path, authoring container only).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this package.
The product (nm_vllm_b200/) never does; it fails loudly when its CUDA extension is missing.
"""
