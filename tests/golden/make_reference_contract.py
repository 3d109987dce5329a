"""Generates tests/golden/reference_contract.json by IMPORTING the reference's own Python modules from
/root/reference (read-only, in this container only; it does not exist on the GPU box):

  * benchmark-vllm/models.py::VLLMConfig parsed from every benchmark-vllm/configs/*.yaml
  * benchmark-vllm/vllm_templates.py::create_vllm_deployment_template / _statefulset_template -> container args
  * VLLMConfig.matches_config / conflicts_with_gpu_resources / get_resource_key on all config pairs
  * benchmark-eval/main.py::process_template on a sample template (function body exec'd: the module needs env/network)
  * benchmark-results/models.py::EvaluationPayload JSON schema + the pk rule of results_manager.py:22

    python tests/golden/make_reference_contract.py
"""
import ast
import importlib.util
import json
import sys
from pathlib import Path

import yaml

REF = Path("/root/reference")
OUT = Path(__file__).parent / "reference_contract.json"


def load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    m = importlib.util.module_from_spec(spec)
    sys.modules[name] = m
    spec.loader.exec_module(m)
    return m


def main():
    models = load(REF / "benchmark-vllm/models.py", "models")            # vllm_templates does `from models import`
    templates = load(REF / "benchmark-vllm/vllm_templates.py", "vllm_templates")
    out = {"configs": {}, "pairs": [], "defaults": models.VLLMConfig(model_name="x").model_dump()}
    cfgs = {}
    for f in sorted((REF / "benchmark-vllm/configs").glob("*.yaml")):
        raw = yaml.safe_load(f.read_text())
        c = models.VLLMConfig(**raw)
        cfgs[f.name] = c
        dep = templates.create_vllm_deployment_template("dep", c, "id-1")
        sts = templates.create_vllm_statefulset_template("dep", c, "id-1")
        out["configs"][f.name] = {
            "yaml": raw,
            "parsed": c.model_dump(),
            "deployment_args": dep["spec"]["template"]["spec"]["containers"][0]["args"],
            "statefulset_args": sts["spec"]["template"]["spec"]["containers"][0]["args"],
            "resource_key": c.get_resource_key(),
        }
    names = sorted(cfgs)
    for a in names:
        for b in names:
            out["pairs"].append([a, b, cfgs[a].matches_config(cfgs[b]), cfgs[a].conflicts_with_gpu_resources(cfgs[b])])

    # process_template: take the function out of benchmark-eval/main.py without importing the FastAPI app
    src = (REF / "benchmark-eval/main.py").read_text()
    fn = next(n for n in ast.parse(src).body if isinstance(n, ast.FunctionDef) and n.name == "process_template")
    ns = {}
    exec(compile(ast.Module([fn], []), "process_template", "exec"), ns)
    template = ("apiVersion: batch/v1\nkind: Job\nmetadata:\n  name: eval-{model_name_lower}\nspec:\n  template:\n"
                "    spec:\n      containers:\n      - name: eval\n        env:\n        - name: MODEL\n"
                "          value: \"{model_name}\"\n        - name: VLLM_MODEL_ENDPOINT\n"
                "          value: \"{inference_engine_url}/v1/completions\"\n")
    cases = [("B200-Llama-3-8B-v1", "http://127.0.0.1:8005"), ("vllm-Qwen2-1-5B-3", "http://vllm-qwen.vllm:8000")]
    out["process_template"] = {"template": template,
                               "cases": [[m, u, ns["process_template"](template, m, u)] for m, u in cases]}

    res = load(REF / "benchmark-results/models.py", "results_models")
    out["evaluation_payload_schema"] = res.EvaluationPayload.model_json_schema()
    p = res.EvaluationPayload(benchmark_name="b", data={"x": 1}, timestamp="2026-01-01T00:00:00Z", model_id="m",
                              tokenizer_id="t", source="s")
    out["evaluation_payload_pk_example"] = {"payload": p.model_dump(),
                                            "pk": f"{p.timestamp}-{p.benchmark_name}-{p.run_id or None}"}
    OUT.write_text(json.dumps(out, indent=1))   # no sort_keys: key order is part of the contract (argv order)
    print("wrote", OUT, OUT.stat().st_size, "bytes")


if __name__ == "__main__":
    main()
