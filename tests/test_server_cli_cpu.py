"""The `vllm serve` command line of the reference's values files, parsed by the drop-in binary
(kge-serve --dry-run needs no GPU).  Flag sets are the ones the chart would pass for each values
file: vllmConfig keys become flags, extraArgs are appended verbatim."""
import json
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "kubernetes-gpu-cluster_b200", "bin", "vllm")

pytestmark = pytest.mark.skipif(not os.path.exists(BIN), reason="server binary not built")


def dry(*argv):
    p = subprocess.run([BIN, *argv, "--dry-run"], capture_output=True, text=True)
    return p.returncode, (json.loads(p.stdout) if p.returncode == 0 else p.stderr)


VALUES_FILES = {
    # ref:values-01-minimal-example.yaml:4-13
    "ex1": (["serve", "facebook/opt-125m", "--host", "0.0.0.0", "--port", "8000"], {"model": "facebook/opt-125m"}),
    # ref:values-01-minimal-example3.yaml:8 (model path is a mounted directory)
    "ex3": (["serve", "/models/Qwen3-0.6B", "--host", "0.0.0.0", "--port", "8000"], {"model": "/models/Qwen3-0.6B"}),
    # ref:values-01-minimal-example6.yaml:16-21
    "ex6": (["serve", "/models/Qwen3-4B", "--host", "0.0.0.0", "--port", "8000", "--gpu-memory-utilization", "0.99",
             "--max-model-len", "128", "--trust-remote-code"],
            {"gpu_memory_utilization": 0.99, "max_model_len": 128, "trust_remote_code": True}),
    # ref:values-01-minimal-example8.yaml:23-38
    "ex8": (["serve", "/models/Qwen2.5-7B", "--host", "0.0.0.0", "--port", "8000", "--max-model-len", "4096",
             "--gpu-memory-utilization", "0.99", "--dtype", "float16", "--disable-custom-all-reduce", "--enforce-eager",
             "--trust-remote-code", "--pipeline-parallel-size", "1", "--tensor-parallel-size", "2"],
            {"max_model_len": 4096, "dtype": "float16", "tensor_parallel_size": 2, "enforce_eager": True,
             "disable_custom_all_reduce": True, "pipeline_parallel_size": 1}),
    # ref:values-01-minimal-example9.yaml:23-39, `=` spelling (commented form at example8.yaml:29)
    "ex9": (["serve", "/models/Qwen3-14B", "--max-model-len=1024", "--dtype=float16", "--tensor-parallel-size=2",
             "--gpu_memory_utilization=0.99"], {"max_model_len": 1024, "tensor_parallel_size": 2, "gpu_memory_utilization": 0.99}),
}


@pytest.mark.parametrize("name", sorted(VALUES_FILES))
def test_values_file_command_lines_parse(name):
    argv, expect = VALUES_FILES[name]
    rc, out = dry(*argv)
    assert rc == 0, out
    for k, v in expect.items():
        assert out[k] == v, (k, out[k], v)
    assert out["port"] == 8000 and out["served_model_name"] == out["model"]


def test_pipeline_parallel_is_refused_with_a_clear_error():
    # ref:values-01-minimal-example4.yaml:16-18 (pipelineParallelSize: 2 + Ray) is out of scope
    rc, err = dry("serve", "/models/Qwen-7B", "--tensor-parallel-size", "1", "--pipeline-parallel-size", "2")
    assert rc == 2 and "pipeline-parallel-size 2 is not supported" in err


def test_bad_command_lines():
    assert dry("serve")[0] == 2                                   # model missing
    assert "unrecognized" in dry("serve", "m", "--no-such-flag")[1]
    assert dry("serve", "m", "--port", "abc")[0] == 2
    assert dry("serve", "m", "--dtype", "int4")[0] == 2
    assert dry("serve", "m", "--gpu-memory-utilization", "1.5")[0] == 2
    rc, out = dry("m", "--served-model-name", "alias", "--load-format", "dummy", "--seed", "7", "--max-num-seqs", "64")
    assert rc == 0 and out["served_model_name"] == "alias" and out["load_format"] == "dummy" and out["seed"] == 7


def test_deploy_values_keep_gpu_count_and_tp_consistent():
    """deploy/values-kge-*.yaml: same schema as the reference's values files (one modelSpec each),
    requestGPU == --tensor-parallel-size, and the flags they pass parse."""
    import glob

    import yaml
    files = sorted(glob.glob(os.path.join(ROOT, "deploy", "values-kge-*.yaml")))
    assert len(files) >= 3
    for f in files:
        spec = yaml.safe_load(open(f))["servingEngineSpec"]
        assert len(spec["modelSpec"]) == 1
        m = spec["modelSpec"][0]
        vc = m.get("vllmConfig", {})
        argv = ["serve", m["modelURL"], "--host", "0.0.0.0", "--port", "8000"]
        if "maxModelLen" in vc:
            argv += ["--max-model-len", str(vc["maxModelLen"])]
        if "gpuMemoryUtilization" in vc:
            argv += ["--gpu-memory-utilization", str(vc["gpuMemoryUtilization"])]
        argv += [str(a) for a in vc.get("extraArgs", [])]
        rc, out = dry(*argv)
        assert rc == 0, (f, out)
        assert out["tensor_parallel_size"] == m["requestGPU"], f


def test_generation_config_defaults_and_extra_eos(tmp_path):
    """vLLM's server takes sampling defaults and extra end-of-sequence ids from the model directory (--generation-config auto):
    Llama-3.1-Instruct stops at <|eot_id|> only because generation_config.json lists it."""
    (tmp_path / "config.json").write_text(json.dumps({"model_type": "llama", "eos_token_id": [128001, 128008, 128009]}))
    (tmp_path / "generation_config.json").write_text(json.dumps({"temperature": 0.6, "top_p": 0.9, "top_k": 20, "do_sample": True,
                                                                  "eos_token_id": [128001, 128008, 128009], "repetition_penalty": 1.05}))
    rc, d = dry("serve", str(tmp_path))
    assert rc == 0 and d["generation_config"] == "auto"
    g = d["generation_defaults"]
    assert (g["temperature"], g["top_p"], g["top_k"]) == (0.6, 0.9, 20) and g["extra_eos_token_ids"] == [128008, 128009]
    assert g["ignored"] == ["repetition_penalty"]
    rc, d = dry("serve", str(tmp_path), "--generation-config", "vllm")
    g = d["generation_defaults"]
    assert (g["temperature"], g["top_p"], g["top_k"]) == (None, None, None) and g["extra_eos_token_ids"] == [128008, 128009]
    other = tmp_path / "gen"
    other.mkdir()
    (other / "generation_config.json").write_text(json.dumps({"temperature": 0.2, "eos_token_id": 7}))
    rc, d = dry("serve", str(tmp_path), f"--generation-config={other}")
    assert d["generation_defaults"]["temperature"] == 0.2 and d["generation_defaults"]["extra_eos_token_ids"] == [7]
    rc, d = dry("serve", "facebook/opt-125m")
    assert d["generation_defaults"] == {"temperature": None, "top_p": None, "top_k": None, "extra_eos_token_ids": [], "ignored": []}
