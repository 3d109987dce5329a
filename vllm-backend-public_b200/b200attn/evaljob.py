"""The evaluation-job side of the reference's loop, without Kubernetes.

benchmark-eval renders a Job template by literal placeholder replacement and hands the YAML to benchmark-deployer,
which creates a Kubernetes Job whose container talks to the engine (reference benchmark-eval/main.py:55-58,108-127;
benchmark-deployer/kubernetes_client.py:133-136).  Here:

  process_template   the same rendering rule (checked against the reference's own function through
                     tests/golden/reference_contract.json)
  render_for_model   what execute_evaluation does before rendering: '_' and '.' in the model name become '-'
                     (main.py:113), and the result must parse as YAML (main.py:116-120)
  run_job_locally    the "kubelet": runs the Job's first container command + args with its env in a subprocess
"""
from __future__ import annotations

import os
import subprocess
import sys
from pathlib import Path
from typing import Dict, List, Optional, Tuple

import yaml

TEMPLATE_PATH = Path(__file__).resolve().parent / "templates" / "evaluate.yaml"


def process_template(template: str, model_name: str, inference_engine_url: str) -> str:
    """{model_name} -> name, {inference_engine_url} -> url, {model_name_lower} -> name.lower(), in that order."""
    return (template.replace("{model_name}", model_name).replace("{inference_engine_url}", inference_engine_url)
            .replace("{model_name_lower}", model_name.lower()))


def render_for_model(model_name: str, inference_engine_url: str, template: Optional[str] = None) -> Tuple[str, dict]:
    """Rendered YAML text and its parsed form for one model / engine URL; raises ValueError on invalid YAML."""
    if template is None:
        template = TEMPLATE_PATH.read_text()
    text = process_template(template, model_name.replace("_", "-").replace(".", "-"), inference_engine_url)
    try:
        doc = yaml.safe_load(text)
    except yaml.YAMLError as e:
        raise ValueError(f"rendered evaluation template is not valid YAML: {e}") from None
    if not isinstance(doc, dict) or doc.get("kind") != "Job":
        raise ValueError("rendered evaluation template is not a Job")
    return text, doc


def job_command(doc: dict) -> Tuple[List[str], Dict[str, str]]:
    """argv and env of the Job's first container."""
    c = doc["spec"]["template"]["spec"]["containers"][0]
    argv = [str(x) for x in (c.get("command") or [])] + [str(x) for x in (c.get("args") or [])]
    env = {e["name"]: str(e.get("value", "")) for e in (c.get("env") or [])}
    if not argv:
        raise ValueError("the Job's container has no command")
    return argv, env


def run_job_locally(doc: dict, extra_args: Optional[List[str]] = None, extra_env: Optional[Dict[str, str]] = None,
                    timeout: Optional[float] = None) -> int:
    """Run the Job's container command here (python -> this interpreter), with the Job's env on top of ours."""
    argv, env = job_command(doc)
    if argv[0] in ("python", "python3"):
        argv[0] = sys.executable
    full_env = dict(os.environ)
    pkg_parent = str(Path(__file__).resolve().parent.parent)
    full_env["PYTHONPATH"] = pkg_parent + os.pathsep + full_env.get("PYTHONPATH", "")
    full_env.update(env)
    full_env.update(extra_env or {})
    return subprocess.call(argv + list(extra_args or []), env=full_env, timeout=timeout)
