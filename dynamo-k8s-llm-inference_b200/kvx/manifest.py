"""Topology from a DynamoGraphDeployment manifest.

The reference's disaggregated manifests are the only place the prefill/decode split is declared
(examples/deploy/vllm/disagg.yaml:18-57, examples/deploy/sglang/disagg.yaml:18-87,
examples/deploy/trtllm/disagg.yaml:18-63, examples/dgdr/trtllm/disagg.yaml).  For every
`spec.services.*` with `componentType: worker` this reads `subComponentType` (or the role flags in
`args`), `replicas`, `resources.limits.gpu`, the model, TP degree and KV page size.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional

import yaml

from .api import HND, NHD

# KV row order each engine's paged cache uses on this path: vLLM's NIXL
# connector forces HND, SGLang pools are token-major, TRT-LLM pools are head-major.
BACKEND_LAYOUT = {"vllm": HND, "sglang": NHD, "trtllm": HND}


@dataclass
class WorkerSpec:
    name: str
    role: str  # "prefill" | "decode" | "aggregated"
    replicas: int
    gpus: int
    tp: int
    page_size: int
    model: Optional[str]
    backend: str
    layout: int
    args: List[str] = field(default_factory=list)
    bootstrap_port: Optional[int] = None   # --disaggregation-bootstrap-port (sglang/disagg.yaml:49-50,84-85)
    transfer_backend: Optional[str] = None  # --disaggregation-transfer-backend (sglang/disagg.yaml:47-48: nixl)


@dataclass
class Topology:
    name: str
    backend: str
    workers: List[WorkerSpec]

    def role(self, role: str) -> List[WorkerSpec]:
        return [w for w in self.workers if w.role == role]

    @property
    def disaggregated(self) -> bool:
        return bool(self.role("prefill")) and bool(self.role("decode"))

    def gpus_needed(self) -> int:
        return sum(w.replicas * w.gpus for w in self.workers if w.role in ("prefill", "decode"))

    def describe(self) -> Dict:
        return {"name": self.name, "backend": self.backend,
                "workers": [{"name": w.name, "role": w.role, "replicas": w.replicas, "gpus": w.gpus, "tp": w.tp,
                             "page_size": w.page_size, "model": w.model, "layout": "HND" if w.layout == HND else "NHD",
                             **({"bootstrap_port": w.bootstrap_port} if w.bootstrap_port else {})}
                            for w in self.workers]}


def _arg_value(args: List[str], *names: str) -> Optional[str]:
    for i, a in enumerate(args):
        for n in names:
            if a == n and i + 1 < len(args):
                return str(args[i + 1])
            if a.startswith(n + "="):
                return a.split("=", 1)[1]
    return None


def _backend_of(container: Dict) -> str:
    text = " ".join(str(x) for x in (container.get("command") or [])) + " " + str(container.get("image", ""))
    for key, name in (("dynamo.vllm", "vllm"), ("dynamo.sglang", "sglang"), ("dynamo.trtllm", "trtllm"),
                      ("vllm-runtime", "vllm"), ("sglang-runtime", "sglang"), ("tensorrtllm-runtime", "trtllm")):
        if key in text:
            return name
    return "vllm"


def _role_of(svc: Dict, args: List[str]) -> str:
    sub = str(svc.get("subComponentType", "")).lower()
    if sub in ("prefill", "decode"):
        return sub
    if "--is-prefill-worker" in args:
        return "prefill"
    if "--is-decode-worker" in args:
        return "decode"
    mode = _arg_value(args, "--disaggregation-mode")
    if mode in ("prefill", "decode"):
        return mode
    return "aggregated"


def parse_manifest(doc: Dict) -> Topology:
    if not isinstance(doc, dict) or doc.get("kind") != "DynamoGraphDeployment":
        raise ValueError("not a DynamoGraphDeployment manifest")
    services = (doc.get("spec") or {}).get("services") or {}
    workers: List[WorkerSpec] = []
    backend = None
    for name, svc in services.items():
        if str(svc.get("componentType", "")).lower() != "worker":
            continue
        container = ((svc.get("extraPodSpec") or {}).get("mainContainer")) or {}
        args = [str(a) for a in (container.get("args") or [])]
        be = _backend_of(container)
        backend = backend or be
        gpus = int(str(((svc.get("resources") or {}).get("limits") or {}).get("gpu", "1")))
        tp = int(_arg_value(args, "--tp", "--tp-size", "--tensor-parallel-size", "-tp") or gpus)
        page = int(_arg_value(args, "--page-size", "--block-size") or 16)
        model = _arg_value(args, "--model", "--model-path", "--served-model-name")
        port = _arg_value(args, "--disaggregation-bootstrap-port")
        workers.append(WorkerSpec(name, _role_of(svc, args), int(svc.get("replicas", 1)), gpus, tp, page, model, be,
                                  BACKEND_LAYOUT[be], args, int(port) if port else None,
                                  _arg_value(args, "--disaggregation-transfer-backend")))
    if not workers:
        raise ValueError("manifest declares no worker services")
    return Topology(str((doc.get("metadata") or {}).get("name", "dgd")), backend or "vllm", workers)


def load_manifest(path: str) -> Topology:
    with open(path) as f:
        docs = [d for d in yaml.safe_load_all(f) if d]
    for d in docs:
        if isinstance(d, dict) and d.get("kind") == "DynamoGraphDeployment":
            return parse_manifest(d)
    raise ValueError(f"{path}: no DynamoGraphDeployment document")


def synthetic_topology(n_prefill: int, n_decode: int, p_tp: int = 1, d_tp: int = 1, backend: str = "vllm",
                       model: Optional[str] = None, decode_layout: Optional[int] = None) -> Topology:
    """The nPmD / TP topologies of BASELINE.json configs #3-#5, in the manifests' vocabulary."""
    lay = BACKEND_LAYOUT[backend]
    return Topology(f"{n_prefill}P{n_decode}D-tp{p_tp}to{d_tp}", backend, [
        WorkerSpec("prefill", "prefill", n_prefill, p_tp, p_tp, 16, model, backend, lay),
        WorkerSpec("decode", "decode", n_decode, d_tp, d_tp, 16, model, backend, lay if decode_layout is None else decode_layout),
    ])
