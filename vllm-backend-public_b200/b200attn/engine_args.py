"""Launch contract of the engine process (SURVEY.md section 8(b) boundary b1).

The reference describes an engine by `VLLMConfig` (reference benchmark-vllm/models.py:5-31) and starts it with the
argument vector built in reference benchmark-vllm/vllm_templates.py:50-87 (StatefulSet) / :221-258 (Deployment), or
through Helm values (reference benchmark-vllm/vllm_manager.py:363-463).  `EngineArgs` carries the same fields under
the same names and defaults, parses that argument vector (every flag the templates emit, including the CPU-era ones
they force on every launch), loads the reference's YAML files (benchmark-vllm/configs/*.yaml) and renders the vector
back, so a config can round-trip:  YAML -> EngineArgs -> argv -> EngineArgs.

Flags that only make sense for the stock vLLM image are accepted and recorded, not obeyed:
  --device cpu                      this engine runs on the B200 it is given; there is no CPU path
  --dtype float32|float16|auto      activations are bf16 (tensor-core GEMMs + the bf16 attention kernels)
  --quantization / --load-format / --max-parallel-loading-workers / --trust-remote-code / --disable-log-*
  --disable-custom-all-reduce       nothing to disable: attention is KV-head sharded without a collective
"""
from __future__ import annotations

import argparse
from dataclasses import asdict, dataclass, field
from pathlib import Path
from typing import Any, Dict, List, Optional

DEFAULT_ADDITIONAL_ARGS = {            # reference benchmark-vllm/models.py:23-28
    "disable-log-stats": True,
    "disable-log-requests": True,
    "enforce-eager": True,
    "disable-custom-all-reduce": True,
}

# flags reference vllm_templates.py:60-66 appends to every launch
FORCED_FLAGS = ["--device", "cpu", "--enforce-eager", "--disable-custom-all-reduce",
                "--max-parallel-loading-workers", "1", "--load-format", "auto",
                "--disable-log-stats", "--disable-log-requests"]


@dataclass
class EngineArgs:
    # --- the fields of VLLMConfig, same names and defaults (reference benchmark-vllm/models.py:7-31) ---
    model_name: str = ""
    gpu_memory_utilization: float = 0.0
    max_num_seqs: int = 2
    block_size: int = 16
    tensor_parallel_size: int = 1
    pipeline_parallel_size: int = 1
    trust_remote_code: bool = False
    dtype: str = "float32"
    max_model_len: Optional[int] = 512
    quantization: Optional[str] = None
    served_model_name: Optional[str] = "test-model-cpu"
    port: int = 8000
    host: str = "0.0.0.0"
    gpu_resource_type: str = "cpu"
    gpu_resource_count: int = 0
    additional_args: Dict[str, Any] = field(default_factory=lambda: dict(DEFAULT_ADDITIONAL_ARGS))
    custom_values_path: Optional[str] = None
    custom_values_content: Optional[str] = None
    # --- flags outside VLLMConfig that the templates force, kept so the vector round-trips ---
    device: Optional[str] = None
    load_format: Optional[str] = None
    max_parallel_loading_workers: Optional[int] = None

    # ------------------------------------------------------------------ reference method mirrors
    def get_resource_key(self) -> str:
        """reference models.py:33-35"""
        return f"{self.gpu_resource_type}:{self.gpu_resource_count}"

    _CORE = ("model_name", "gpu_memory_utilization", "max_num_seqs", "block_size", "tensor_parallel_size",
             "pipeline_parallel_size", "trust_remote_code", "dtype", "max_model_len", "quantization",
             "served_model_name", "gpu_resource_type", "gpu_resource_count")

    def matches_config(self, other: "EngineArgs") -> bool:
        """Deployment reuse test, reference models.py:37-55: every core field and the additional args agree."""
        return all(getattr(self, f) == getattr(other, f) for f in self._CORE) and \
            self.additional_args == other.additional_args

    def conflicts_with_gpu_resources(self, other: "EngineArgs") -> bool:
        """reference models.py:57-72: the same resource type conflicts; so do two MIG resources of one slice type."""
        if self.gpu_resource_type == other.gpu_resource_type:
            return True
        mig = "nvidia.com/mig-"
        if self.gpu_resource_type.startswith(mig) and other.gpu_resource_type.startswith(mig):
            return self.gpu_resource_type.split(mig)[-1] == other.gpu_resource_type.split(mig)[-1]
        return False

    # ------------------------------------------------------------------ constructors
    @classmethod
    def from_config_dict(cls, d: Dict[str, Any]) -> "EngineArgs":
        """The YAML / JSON shape of VLLMConfig (reference benchmark-vllm/configs/*.yaml)."""
        known = {f for f in cls.__dataclass_fields__}
        unknown = set(d) - known
        if unknown:
            raise ValueError(f"unknown VLLMConfig field(s): {sorted(unknown)}")
        if "model_name" not in d:
            raise ValueError("model_name is required")
        a = cls(**{k: v for k, v in d.items()})
        if d.get("additional_args") is None:
            a.additional_args = dict(DEFAULT_ADDITIONAL_ARGS) if "additional_args" not in d else {}
        return a

    @classmethod
    def from_yaml(cls, path) -> "EngineArgs":
        import yaml
        return cls.from_config_dict(yaml.safe_load(Path(path).read_text()))

    @classmethod
    def from_cli(cls, argv: List[str]) -> "EngineArgs":
        """Parse the argument vector the reference renders (vllm_templates.py:50-87).  Unknown `--key [value]` pairs
        land in additional_args the way the templates emit them (bare flag = True)."""
        p = argparse.ArgumentParser(prog="b200attn-server", add_help=True, allow_abbrev=False)
        p.add_argument("--model", required=True)
        p.add_argument("--gpu-memory-utilization", type=float, default=0.9)
        p.add_argument("--max-num-seqs", type=int, default=64)
        p.add_argument("--block-size", type=int, default=16)
        p.add_argument("--tensor-parallel-size", type=int, default=1)
        p.add_argument("--pipeline-parallel-size", type=int, default=1)
        p.add_argument("--trust-remote-code", action="store_true")
        p.add_argument("--dtype", default="auto")
        p.add_argument("--max-model-len", type=int, default=None)
        p.add_argument("--quantization", default=None)
        p.add_argument("--served-model-name", default=None)
        p.add_argument("--port", type=int, default=8000)
        p.add_argument("--host", default="0.0.0.0")
        p.add_argument("--device", default=None)
        p.add_argument("--load-format", default=None)
        p.add_argument("--max-parallel-loading-workers", type=int, default=None)
        ns, rest = p.parse_known_args(argv)
        extra: Dict[str, Any] = {}
        i = 0
        while i < len(rest):
            tok = rest[i]
            if not tok.startswith("--"):
                raise ValueError(f"unexpected argument {tok!r}")
            key = tok[2:]
            if "=" in key:
                key, val = key.split("=", 1)
                extra[key] = _coerce(val)
                i += 1
            elif i + 1 < len(rest) and not rest[i + 1].startswith("--"):
                extra[key] = _coerce(rest[i + 1])
                i += 2
            else:
                extra[key] = True
                i += 1
        return cls(model_name=ns.model, gpu_memory_utilization=ns.gpu_memory_utilization, max_num_seqs=ns.max_num_seqs,
                   block_size=ns.block_size, tensor_parallel_size=ns.tensor_parallel_size,
                   pipeline_parallel_size=ns.pipeline_parallel_size, trust_remote_code=ns.trust_remote_code,
                   dtype=ns.dtype, max_model_len=ns.max_model_len, quantization=ns.quantization,
                   served_model_name=ns.served_model_name, port=ns.port, host=ns.host, additional_args=extra,
                   device=ns.device, load_format=ns.load_format,
                   max_parallel_loading_workers=ns.max_parallel_loading_workers)

    # ------------------------------------------------------------------ rendering
    def to_cli_args(self, forced: bool = True) -> List[str]:
        """The vector reference vllm_templates.py:50-87 builds for this config (`forced` = with the flags it
        appends to every launch)."""
        a = ["--model", self.model_name,
             "--gpu-memory-utilization", str(self.gpu_memory_utilization),
             "--max-num-seqs", str(self.max_num_seqs),
             "--block-size", str(self.block_size),
             "--tensor-parallel-size", str(self.tensor_parallel_size),
             "--pipeline-parallel-size", str(self.pipeline_parallel_size),
             "--dtype", self.dtype,
             "--port", str(self.port),
             "--host", self.host]
        if forced:
            a += FORCED_FLAGS
        if self.trust_remote_code:
            a.append("--trust-remote-code")
        if self.max_model_len:
            a += ["--max-model-len", str(self.max_model_len)]
        if self.quantization:
            a += ["--quantization", self.quantization]
        if self.served_model_name:
            a += ["--served-model-name", self.served_model_name]
        for k, v in (self.additional_args or {}).items():
            if isinstance(v, bool) and v:
                a.append(f"--{k}")
            elif not isinstance(v, bool):
                a += [f"--{k}", str(v)]
        return a

    def to_config_dict(self) -> Dict[str, Any]:
        d = asdict(self)
        for k in ("device", "load_format", "max_parallel_loading_workers"):
            d.pop(k)
        return d

    # ------------------------------------------------------------------ what the engine does with them
    def extra(self, key: str, default=None):
        return (self.additional_args or {}).get(key, default)

    def engine_kwargs(self) -> Dict[str, Any]:
        """Arguments of b200attn.engine.LLMEngine, with the notes of the module docstring applied."""
        if self.pipeline_parallel_size != 1:
            raise ValueError("pipeline_parallel_size > 1 is not supported (one process per GPU, KV-head sharding only)")
        if self.tensor_parallel_size not in (1, 2, 4, 8):
            raise ValueError(f"tensor_parallel_size {self.tensor_parallel_size}: one process per GPU of one node, "
                             "1, 2, 4 or 8 (b200attn.tp)")
        if self.block_size not in (8, 16, 32):
            raise ValueError(f"block_size {self.block_size}: the kernels are built for 8, 16 and 32")
        kw = dict(max_num_seqs=self.max_num_seqs, block_size=self.block_size,
                  gpu_memory_utilization=self.gpu_memory_utilization,
                  max_model_len=self.max_model_len or 0,
                  enforce_eager=False,     # --enforce-eager is a CPU-era flag of the templates (vllm_templates.py:61)
                  kv_cache_dtype=str(self.extra("kv-cache-dtype", "auto")),
                  seed=int(self.extra("seed", 0)))
        if self.extra("max-num-batched-tokens") is not None:
            kw["max_num_batched_tokens"] = int(self.extra("max-num-batched-tokens"))
        if self.extra("num-gpu-blocks-override") is not None:
            kw["num_blocks"] = int(self.extra("num-gpu-blocks-override"))
        if self.extra("b200-eager"):
            kw["enforce_eager"] = True
        return kw

    @property
    def public_name(self) -> str:
        return self.served_model_name or self.model_name


def _coerce(v: str):
    for cast in (int, float):
        try:
            return cast(v)
        except ValueError:
            pass
    if v.lower() in ("true", "false"):
        return v.lower() == "true"
    return v
