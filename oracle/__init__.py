"""CPU oracle for the serving-engine hot path.  TEST INFRASTRUCTURE ONLY.

PARITY UNPINNED: the reference (/root/reference) holds no tests, golden vectors or
compute code for this path.  The arithmetic lives in a third-party
dependency, vLLM, pinned by container tag v0.8.5 / v0.8.4
(ref:values-01-minimal-example.yaml:5-6, ref:values-01-minimal-example2.yaml:6-7).
This package restates vLLM's published op order and cast points in plain PyTorch on
the CPU (citations are to the vLLM 0.22.0 sources installed in the build container,
prefix ``vllm:``) and is anchored by fixtures generated in the build container from
vLLM's own importable pure-PyTorch functions and from HF transformers
(tests/golden/make_golden.py), from step-by-step traces of vLLM's real v1 Scheduler without and
with prefix caching (make_sched_golden.py, make_sched_prefix_golden.py), and from tokens and
log-probabilities of the installed vLLM 0.22.0 running its own CUDA kernels on a B200
(make_vllm_cuda_golden.py).  What stays unpinned is the exact version the reference names
(v0.8.x) and vLLM's CPU backend: neither can run here.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this package, and only as the checker or the timed CPU baseline.
The product (kubernetes-gpu-cluster_b200/) never imports it.
"""
