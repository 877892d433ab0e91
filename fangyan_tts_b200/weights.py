"""Weight loading for the drop-in CLIs: `--onnx_path` -> named fp32 tensors for fy_load_campplus / fy_load_s3tok.

The reference hands `--onnx_path` to onnxruntime.InferenceSession (tools/extract_embedding.py:94,
tools/extract_speech_token.py:90).  Neither `onnx` nor `onnxruntime` is installable offline, so this module reads
the ONNX protobuf directly (a small wire-format reader: initialisers, Constant nodes and the node list with the
attributes that matter) and also accepts a torch state dict (`.pt` / `.pth`) or an `.npz` with the tensor names of
synth_weights.py.

An exported graph does not carry PyTorch module names: the exporter renames Linear weights `onnx::MatMul_N` and
stores them [in, out], folds a BatchNorm that follows a convolution into that convolution (weight renamed
`onnx::Conv_N`, a bias appears, the running statistics disappear) and keeps the names of whatever it does not touch
-- and those are the names of the ORIGINAL project's modules, not ours.  So an `.onnx` file is mapped by TOPOLOGY,
not by name: the nodes that consume weights are walked in graph order and matched, by operator type and full weight
shape, against the sequence the hyper-parameter table (configs.py) implies.  A graph whose structure differs fails
with a message naming the first unit that does not fit and what was found instead; nothing is guessed.

UNVERIFIED against the real campplus.onnx / speech_tokenizer_v3.onnx (not available offline): the walk is tested on
graphs written by tests/onnx_writer.py with exporter conventions (tests/test_weights_cpu.py).
"""
from __future__ import annotations

import struct
from dataclasses import dataclass, field
from pathlib import Path
from typing import Dict, List, Optional, Tuple

import numpy as np

_ONNX_DTYPES = {1: np.float32, 10: np.float16, 11: np.float64, 6: np.int32, 7: np.int64, 9: np.bool_}


class OnnxStructureError(ValueError):
    """The graph does not have the structure the hyper-parameter table describes."""


# ------------------------------------------------------------------------------------------ protobuf wire format
def _varint(buf: memoryview, pos: int):
    val, shift = 0, 0
    while True:
        b = buf[pos]
        pos += 1
        val |= (b & 0x7F) << shift
        if not b & 0x80:
            return val, pos
        shift += 7


def _fields(buf: memoryview):
    """Yield (field_number, wire_type, value) over one protobuf message."""
    pos, n = 0, len(buf)
    while pos < n:
        key, pos = _varint(buf, pos)
        fnum, wt = key >> 3, key & 7
        if wt == 0:
            v, pos = _varint(buf, pos)
        elif wt == 1:
            v = buf[pos:pos + 8]
            pos += 8
        elif wt == 2:
            ln, pos = _varint(buf, pos)
            v = buf[pos:pos + ln]
            pos += ln
        elif wt == 5:
            v = buf[pos:pos + 4]
            pos += 4
        else:
            raise ValueError(f"unsupported protobuf wire type {wt}")
        yield fnum, wt, v


def _packed_varints(v: memoryview) -> List[int]:
    vals, p = [], 0
    while p < len(v):
        d, p = _varint(v, p)
        vals.append(d)
    return vals


def _signed(v: int) -> int:
    return v - (1 << 64) if v >= 1 << 63 else v


def _tensor(buf: memoryview):
    """onnx.TensorProto -> (name, ndarray) for the dtypes initialisers use."""
    dims, dtype, name, raw, floats, int64s = [], 1, "", None, None, None
    for f, wt, v in _fields(buf):
        if f == 1:      # dims (possibly packed)
            dims.extend([v] if wt == 0 else _packed_varints(v))
        elif f == 2:
            dtype = v
        elif f == 8:
            name = bytes(v).decode()
        elif f == 9:
            raw = v
        elif f == 4:    # float_data, packed
            floats = np.frombuffer(v, dtype="<f4") if wt == 2 else np.array([struct.unpack("<f", v)[0]], np.float32)
        elif f == 7:    # int64_data, packed varints
            int64s = np.array([_signed(x) for x in (_packed_varints(v) if wt == 2 else [v])], dtype=np.int64)
    if dtype not in _ONNX_DTYPES:
        return name, None
    if raw is not None:
        arr = np.frombuffer(raw, dtype=np.dtype(_ONNX_DTYPES[dtype]).newbyteorder("<"))
    elif floats is not None:
        arr = floats
    elif int64s is not None:
        arr = int64s
    else:
        arr = np.zeros(0, _ONNX_DTYPES[dtype])
    return name, arr.reshape(dims) if dims else arr


@dataclass
class Node:
    op: str
    name: str = ""
    inputs: List[str] = field(default_factory=list)
    outputs: List[str] = field(default_factory=list)
    attrs: Dict[str, object] = field(default_factory=dict)


def _attribute(buf: memoryview):
    """onnx.AttributeProto -> (name, python value) for float / int / ints / floats / tensor attributes."""
    name, val = "", None
    ints, floats = [], []
    for f, wt, v in _fields(buf):
        if f == 1:
            name = bytes(v).decode()
        elif f == 2:
            val = struct.unpack("<f", v)[0]
        elif f == 3:
            val = _signed(v)
        elif f == 4:
            val = bytes(v)
        elif f == 5:
            val = _tensor(v)[1]
        elif f == 7:
            floats.extend(np.frombuffer(v, dtype="<f4").tolist() if wt == 2 else [struct.unpack("<f", v)[0]])
        elif f == 8:
            ints.extend([_signed(x) for x in _packed_varints(v)] if wt == 2 else [_signed(v)])
    if ints:
        val = ints
    elif floats:
        val = floats
    return name, val


def _node(buf: memoryview) -> Node:
    n = Node(op="")
    for f, wt, v in _fields(buf):
        if f == 1:
            n.inputs.append(bytes(v).decode())
        elif f == 2:
            n.outputs.append(bytes(v).decode())
        elif f == 3:
            n.name = bytes(v).decode()
        elif f == 4:
            n.op = bytes(v).decode()
        elif f == 5:
            k, a = _attribute(v)
            n.attrs[k] = a
    return n


def read_onnx_graph(path) -> Tuple[Dict[str, np.ndarray], List[Node]]:
    """ModelProto.graph (field 7) -> (initialisers incl. Constant-node values and Identity aliases, nodes in order)."""
    data = memoryview(Path(path).read_bytes())
    inits: Dict[str, np.ndarray] = {}
    nodes: List[Node] = []
    for f, wt, v in _fields(data):
        if f == 7 and wt == 2:
            for gf, gwt, gv in _fields(v):
                if gf == 5 and gwt == 2:
                    name, arr = _tensor(gv)
                    if arr is not None:
                        inits[name] = arr
                elif gf == 1 and gwt == 2:
                    nodes.append(_node(gv))
    if not inits and not nodes:
        raise ValueError(f"{path}: no graph found (not an ONNX model?)")
    for n in nodes:                                   # values the exporter left as Constant nodes / shared via Identity
        if n.op == "Constant" and n.outputs and isinstance(n.attrs.get("value"), np.ndarray):
            inits[n.outputs[0]] = n.attrs["value"]
        elif n.op == "Identity" and n.inputs and n.inputs[0] in inits and n.outputs:
            inits[n.outputs[0]] = inits[n.inputs[0]]
    return inits, nodes


def read_onnx_initializers(path) -> Dict[str, np.ndarray]:
    inits, _ = read_onnx_graph(path)
    if not inits:
        raise ValueError(f"{path}: no initialisers found (not an ONNX model?)")
    return inits


# ------------------------------------------------------------------------------------------ topology walk
@dataclass
class _W:
    """One weight-consuming node, reduced to what the mapping needs."""
    op: str                       # Conv | Linear | BatchNorm | LayerNorm
    w: Optional[np.ndarray]       # Conv [Cout, Cin/g, k...]; Linear [out, in] (already transposed); None for norms
    b: Optional[np.ndarray]
    out: str
    extra: Dict[str, object]
    where: str                    # node name / output, for messages

    @property
    def sig(self) -> str:
        if self.op in ("BatchNorm", "LayerNorm"):
            return f"{self.op}[{self.extra['c']}]"
        return f"{self.op}{list(self.w.shape)}{'+bias' if self.b is not None else ''}"


def _weighted_nodes(inits: Dict[str, np.ndarray], nodes: List[Node]) -> List[_W]:
    """Graph order list of Conv / MatMul / Gemm / BatchNormalization / LayerNormalization nodes whose parameters are
    constants.  A MatMul's bias is the constant operand of the Add that consumes its output; MatMul stores [in, out]
    (Gemm too unless transB), both are returned [out, in]."""
    consumers: Dict[str, List[Node]] = {}
    for n in nodes:
        for i in n.inputs:
            consumers.setdefault(i, []).append(n)
    fl = lambda a: a is not None and a.dtype.kind == "f"
    out: List[_W] = []
    for n in nodes:
        get = lambda k: inits.get(n.inputs[k]) if len(n.inputs) > k and n.inputs[k] else None
        where = n.name or (n.outputs[0] if n.outputs else n.op)
        if n.op == "Conv" and fl(get(1)):
            out.append(_W("Conv", np.asarray(get(1), np.float32), None if get(2) is None else np.asarray(get(2), np.float32),
                          n.outputs[0], {"group": int(n.attrs.get("group", 1) or 1), "strides": n.attrs.get("strides"),
                                         "dilations": n.attrs.get("dilations")}, where))
        elif n.op in ("MatMul", "Gemm") and fl(get(1)) and get(1).ndim == 2:
            w = np.asarray(get(1), np.float32)
            if not (n.op == "Gemm" and int(n.attrs.get("transB", 0) or 0)):
                w = np.ascontiguousarray(w.T)
            b = np.asarray(get(2), np.float32) if n.op == "Gemm" and fl(get(2)) else None
            res = n.outputs[0]
            if b is None:
                for c in consumers.get(n.outputs[0], []):
                    if c.op == "Add":
                        other = [inits.get(i) for i in c.inputs if i != n.outputs[0]]
                        if other and fl(other[0]) and other[0].size == w.shape[0]:
                            b, res = np.asarray(other[0], np.float32).reshape(-1), c.outputs[0]
                            break
            out.append(_W("Linear", w, b, res, {}, where))
        elif n.op == "BatchNormalization" and all(fl(get(k)) for k in (1, 2, 3, 4)):
            out.append(_W("BatchNorm", None, None, n.outputs[0],
                          {"c": int(get(1).size), "scale": get(1), "bias": get(2), "mean": get(3), "var": get(4),
                           "eps": float(n.attrs.get("epsilon", 1e-5) or 1e-5), "input": n.inputs[0]}, where))
        elif n.op == "LayerNormalization" and fl(get(1)):
            out.append(_W("LayerNorm", None, None, n.outputs[0],
                          {"c": int(get(1).size), "scale": get(1), "bias": get(2) if fl(get(2)) else np.zeros(get(1).size, np.float32),
                           "eps": float(n.attrs.get("epsilon", 1e-5) or 1e-5)}, where))
    return out


class _Walk:
    def __init__(self, items: List[_W], path):
        self.items, self.i, self.path = items, 0, path

    def fail(self, unit: str, want: str):
        got = self.items[self.i].sig + f" at node '{self.items[self.i].where}'" if self.i < len(self.items) else "end of graph"
        ctx = ", ".join(w.sig for w in self.items[max(0, self.i - 2):self.i + 3])
        raise OnnxStructureError(
            f"{self.path}: graph structure differs from the hyper-parameter table (configs.py) at unit '{unit}': expected "
            f"{want}, found {got} (weighted node {self.i} of {len(self.items)}; neighbourhood: {ctx}).  If the real model "
            "differs from the published architecture, correct the table (FY_CAMPPLUS_CONFIG / FY_S3TOK_CONFIG) rather "
            "than the weights.")

    def peek(self) -> Optional[_W]:
        return self.items[self.i] if self.i < len(self.items) else None

    def take(self, unit: str, op: str, shape=None, c=None) -> _W:
        w = self.peek()
        want = f"{op}{list(shape)}" if shape is not None else f"{op}[{c}]"
        if w is None or w.op != op:
            self.fail(unit, want)
        if shape is not None and tuple(w.w.shape) != tuple(shape):
            self.fail(unit, want)
        if c is not None and w.extra["c"] != c:
            self.fail(unit, want)
        self.i += 1
        return w

    def done(self, unit="end of network"):
        if self.i != len(self.items):
            self.fail(unit, "no further weighted nodes")


def _put_bn(state, name, bn: Optional[_W], c: int, folded_bias: Optional[np.ndarray], affine: bool, cfg_eps: float):
    """Our loaders fold `name.{running_mean,running_var,weight,bias}` with cfg_eps.  A real BatchNormalization node is
    re-expressed with that epsilon (var' = var + eps_node - eps_cfg); a BatchNorm the exporter already folded into the
    preceding convolution becomes the identity carrying the convolution's bias."""
    if bn is not None:
        mean, var = np.asarray(bn.extra["mean"], np.float32), np.asarray(bn.extra["var"], np.float32) + np.float32(bn.extra["eps"] - cfg_eps)
        scale, bias = np.asarray(bn.extra["scale"], np.float32), np.asarray(bn.extra["bias"], np.float32)
    else:
        mean, var = np.zeros(c, np.float32), np.full(c, 1.0 - cfg_eps, np.float32)
        scale, bias = np.ones(c, np.float32), (np.zeros(c, np.float32) if folded_bias is None else folded_bias.astype(np.float32))
    state[name + ".running_mean"], state[name + ".running_var"] = mean, var
    if affine:
        state[name + ".weight"], state[name + ".bias"] = scale, bias
    elif not (np.allclose(scale, 1.0) and np.allclose(bias, 0.0)):
        # a non-affine BN that arrives with affine terms (folded form): express through mean/var instead
        inv = scale / np.sqrt(var + np.float32(cfg_eps))
        state[name + ".running_var"] = (1.0 / np.maximum(inv, 1e-30) ** 2 - cfg_eps).astype(np.float32)
        state[name + ".running_mean"] = (mean - bias / np.maximum(inv, 1e-30)).astype(np.float32)


def campplus_from_onnx(path, cfg) -> Dict[str, np.ndarray]:
    """Map campplus.onnx onto the tensor names fy_load_campplus expects, by graph order and shape (module docstring)."""
    inits, nodes = read_onnx_graph(path)
    wk = _Walk(_weighted_nodes(inits, nodes), path)
    st: Dict[str, np.ndarray] = {}
    eps = cfg.bn_eps

    def conv_then_bn(unit, conv_name, bn_name, shape, affine=True, linear_ok=False):
        """Conv (or, for 1x1 convolutions on pooled vectors, a MatMul/Gemm) followed by a BatchNorm that is either its
        own node or already folded into the convolution (bias present, no statistics)."""
        w = wk.peek()
        if linear_ok and w is not None and w.op == "Linear":
            w = wk.take(unit, "Linear", shape[:2])
            st[conv_name + ".weight"] = w.w.reshape(shape)
        else:
            w = wk.take(unit, "Conv", shape)
            st[conv_name + ".weight"] = w.w
        nxt = wk.peek()
        if nxt is not None and nxt.op == "BatchNorm" and nxt.extra["c"] == shape[0] and nxt.extra["input"] == w.out:
            if w.b is not None:
                wk.fail(unit, "a bias-free convolution in front of its BatchNormalization node")
            _put_bn(st, bn_name, wk.take(unit, "BatchNorm", c=shape[0]), shape[0], None, affine, eps)
        else:
            if w.b is None:
                wk.fail(unit, f"BatchNorm[{shape[0]}] after {conv_name} (or a convolution bias if the exporter folded it)")
            _put_bn(st, bn_name, None, shape[0], w.b, affine, eps)

    def bn_then_conv(unit, bn_name, c_in, conv_name, shape, post_bn=None, bias=False):
        _put_bn(st, bn_name, wk.take(unit, "BatchNorm", c=c_in), c_in, None, True, eps)
        if post_bn:
            conv_then_bn(unit, conv_name, post_bn, shape)
        else:
            w = wk.take(unit, "Conv", shape)
            st[conv_name + ".weight"] = w.w
            if bias:
                if w.b is None:
                    wk.fail(unit, f"a bias on {conv_name}")
                st[conv_name + ".bias"] = w.b
            elif w.b is not None and np.abs(w.b).max() > 0:
                wk.fail(unit, f"no bias on {conv_name}")

    c = cfg.fcm_channels
    conv_then_bn("head.conv1", "head.conv1", "head.bn1", (c, 1, 3, 3))
    i = 0
    for nb in cfg.fcm_blocks:
        for j in range(nb):
            pre = f"head.blocks.{i}"
            conv_then_bn(pre + ".conv1", pre + ".conv1", pre + ".bn1", (c, c, 3, 3))
            conv_then_bn(pre + ".conv2", pre + ".conv2", pre + ".bn2", (c, c, 3, 3))
            if j == 0:
                conv_then_bn(pre + ".shortcut", pre + ".sc_conv", pre + ".sc_bn", (c, c, 1, 1))
            i += 1
    conv_then_bn("head.conv2", "head.conv2", "head.bn2", (c, c, 3, 3))
    ch = c * (cfg.feat_dim // 8)
    conv_then_bn("tdnn", "tdnn", "tdnn_nl.0", (cfg.init_channels, ch, 5))
    ch, bnc = cfg.init_channels, cfg.bn_size * cfg.growth
    for b, (nl, k, _d) in enumerate(zip(cfg.block_layers, cfg.block_kernel, cfg.block_dilation)):
        for l in range(nl):
            pre, cin = f"blocks.{b}.{l}", ch + l * cfg.growth
            bn_then_conv(pre, pre + ".nl1.0", cin, pre + ".lin1", (bnc, cin, 1), post_bn=pre + ".nl2.0")
            for nm, shape, bias in ((".cam.local", (cfg.growth, bnc, k), False), (".cam.lin1", (bnc // 2, bnc, 1), True),
                                    (".cam.lin2", (cfg.growth, bnc // 2, 1), True)):
                w = wk.take(pre + nm, "Conv", shape)
                st[pre + nm + ".weight"] = w.w
                if bias:
                    if w.b is None:
                        wk.fail(pre + nm, "a bias")
                    st[pre + nm + ".bias"] = w.b
        ch += nl * cfg.growth
        bn_then_conv(f"transits.{b}", f"transits.{b}.nl.0", ch, f"transits.{b}.lin", (ch // 2, ch, 1))
        ch //= 2
    _put_bn(st, "out_nl.0", wk.take("out_nl", "BatchNorm", c=ch), ch, None, True, eps)
    conv_then_bn("dense", "dense", "dense_bn", (cfg.embed_dim, 2 * ch, 1), affine=False, linear_ok=True)
    wk.done()
    return st


def s3tok_from_onnx(path, cfg) -> Dict[str, np.ndarray]:
    """Map speech_tokenizer_v3.onnx onto the names fy_load_s3tok expects.  Linear layers arrive as MatMul [in, out]
    (+ Add); q / k / v / out are told apart by graph order and by the key projection having no bias; attention's own
    MatMuls (no constant operand) and the rotary / FSQ constants are not weighted nodes and are skipped."""
    inits, nodes = read_onnx_graph(path)
    items = [w for w in _weighted_nodes(inits, nodes)]
    wk = _Walk(items, path)
    st: Dict[str, np.ndarray] = {}
    d, f = cfg.d_model, cfg.d_ff

    def conv(unit, name, shape, bias=True):
        w = wk.take(unit, "Conv", shape)
        st[name + ".weight"] = w.w
        if bias:
            if w.b is None:
                wk.fail(unit, "a bias")
            st[name + ".bias"] = w.b

    def linear(unit, name, n_out, n_in, bias=True):
        w = wk.take(unit, "Linear", (n_out, n_in))
        if (w.b is not None) != bias:
            wk.fail(unit, f"Linear[{n_out}, {n_in}] {'with' if bias else 'WITHOUT'} bias ({name})")
        st[name + ".weight"] = w.w
        if bias:
            st[name + ".bias"] = w.b

    def lnorm(unit, name):
        w = wk.take(unit, "LayerNorm", c=d)
        want = cfg.ln_eps if name.endswith("attn_ln") or cfg.mlp_ln_eps is None else cfg.mlp_ln_eps
        if abs(w.extra["eps"] - want) > 1e-12:
            raise OnnxStructureError(f"{path}: {unit} has epsilon {w.extra['eps']:g}, the hyper-parameter table says "
                                     f"{want:g} (set ln_eps / mlp_ln_eps through FY_S3TOK_CONFIG)")
        st[name + ".weight"], st[name + ".bias"] = np.asarray(w.extra["scale"], np.float32), np.asarray(w.extra["bias"], np.float32)

    conv("conv1", "conv1", (d, cfg.n_mels, 3))
    conv("conv2", "conv2", (d, d, 3))
    for i in range(cfg.n_layer):
        p = f"blocks.{i}."
        lnorm(p + "attn_ln", p + "attn_ln")
        linear(p + "attn.query", p + "attn.query", d, d)
        linear(p + "attn.key", p + "attn.key", d, d, bias=False)
        linear(p + "attn.value", p + "attn.value", d, d)
        # the FSMN branch (depthwise Conv over V) and the output projection follow in either order
        for _ in range(2):
            nxt = wk.peek()
            if nxt is not None and nxt.op == "Conv" and p + "attn.fsmn.weight" not in st:
                w = wk.take(p + "attn.fsmn", "Conv", (d, 1, cfg.fsmn_kernel))
                if w.extra["group"] != d:
                    wk.fail(p + "attn.fsmn", f"a depthwise convolution (group {d})")
                st[p + "attn.fsmn.weight"] = w.w
            elif p + "attn.out.weight" not in st:
                linear(p + "attn.out", p + "attn.out", d, d)
        if p + "attn.fsmn.weight" not in st or p + "attn.out.weight" not in st:
            wk.fail(p + "attn", "the FSMN depthwise Conv and the output projection")
        lnorm(p + "mlp_ln", p + "mlp_ln")
        linear(p + "fc1", p + "fc1", f, d)
        linear(p + "fc2", p + "fc2", d, f)
    linear("project_down", "project_down", cfg.fsq_dim, d)
    wk.done()
    return st


# ------------------------------------------------------------------------------------------ entry points
def load_state(path, kind: Optional[str] = None, cfg=None) -> Dict[str, np.ndarray]:
    """Named tensors from .onnx / .pt / .pth / .npz.  kind = 'campplus' | 's3tok' selects the topology mapping for
    `.onnx` files (cfg = the hyper-parameter table); without it the raw initialisers are returned."""
    p = Path(path)
    suf = p.suffix.lower()
    if suf == ".onnx":
        if kind == "campplus":
            from .configs import campplus_config
            return campplus_from_onnx(p, cfg or campplus_config())
        if kind == "s3tok":
            from .configs import s3_config
            return s3tok_from_onnx(p, cfg or s3_config())
        return read_onnx_initializers(p)
    if suf == ".npz":
        with np.load(p) as z:
            return {k: z[k] for k in z.files}
    if suf in (".pt", ".pth", ".bin"):
        import torch
        sd = torch.load(p, map_location="cpu", weights_only=True)
        if isinstance(sd, dict) and "state_dict" in sd:
            sd = sd["state_dict"]
        return {k: v.detach().cpu().numpy() for k, v in sd.items()}
    raise ValueError(f"{path}: unknown weight file type '{suf}' (expected .onnx, .pt or .npz)")


def expected_shapes(kind: str, cfg) -> Dict[str, Tuple[int, ...]]:
    """Full shapes of every tensor the C loaders look up (they check element counts; this catches a transposed square
    matrix or a mis-ordered convolution kernel before it reaches them)."""
    from . import synth_weights
    fn = synth_weights.campplus_shapes if kind == "campplus" else synth_weights.s3tok_shapes
    return fn(cfg)


def check_shapes(state: Dict[str, np.ndarray], kind: str, cfg) -> None:
    bad, missing = [], []
    for name, shape in expected_shapes(kind, cfg).items():
        if name not in state:
            missing.append(name)
        elif tuple(np.shape(state[name])) != tuple(shape):
            bad.append(f"{name}: {tuple(np.shape(state[name]))} != {tuple(shape)}")
    if missing:
        raise KeyError(f"{kind} weights lack {len(missing)} tensors, e.g. {missing[:5]}")
    if bad:
        raise ValueError(f"{kind} weights have wrong shapes: {bad[:5]}")


def as_torch_state(state: Dict[str, np.ndarray], expected_names):
    """Check the names the loader needs are present and convert to torch tensors."""
    import torch
    missing = [n for n in expected_names if n not in state]
    if missing:
        raise KeyError(f"weight file lacks {len(missing)} of {len(expected_names)} expected tensors, e.g. {missing[:5]}")
    return {n: torch.from_numpy(np.ascontiguousarray(state[n], dtype=np.float32)) for n in expected_names}
