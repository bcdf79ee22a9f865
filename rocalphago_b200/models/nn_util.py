"""NeuralNetBase and friends on the B200 path (drop-in for AlphaGo/models/nn_util.py).

`NeuralNetBase(feature_list, **kwargs)` keeps the reference's surface -- `.preprocessor`,
`.model`, `.forward(ndarray)`, `load_model` / `save_model`, the `@neuralnet` registry and a
`Bias` name -- but `.model` is a `Network` (layer list + weights + a handle on the CUDA
forward), not a Keras object: training is out of scope.
"""
import ctypes
import json
import os

import numpy as np
import torch

from .. import _cabi, hdf5_lite
from ..preprocessing import Preprocess, DTYPE_BF16, DTYPE_FP16, DTYPE_CODES, LAYOUT_CONV_IN, analyze_into
from ..state import _ptr, _stream, _require_cuda

HEAD_POLICY, HEAD_VALUE = 0, 1
MAX_LAYERS = 32


class _NetDesc(ctypes.Structure):
    _fields_ = [("board", ctypes.c_int), ("in_planes", ctypes.c_int), ("layers", ctypes.c_int),
                ("filters", ctypes.c_int * MAX_LAYERS), ("widths", ctypes.c_int * MAX_LAYERS),
                ("head", ctypes.c_int), ("dense_units", ctypes.c_int), ("dtype", ctypes.c_int),
                ("kernel_flip", ctypes.c_int), ("resnet", ctypes.c_int), ("unit_end", ctypes.c_int * MAX_LAYERS)]


class Bias(object):
    """Name kept for `custom_objects={'Bias': Bias}` call sites (nn_util.py:118-133).  On this
    path the per-location bias is the last array of a policy Network's weights."""


class Network(object):
    """What sits behind `.model`: the layer stack of policy.py:94-138 / value.py:14-28.

    specs: list of (filters, filter_width) for the k x k ReLU layers; a 1x1 convolution to one
    plane follows, then Flatten + Bias + softmax (head='policy') or Flatten + Dense(dense_units)
    + Dense(1, tanh) (head='value').  Weights are kept as numpy arrays in Keras'
    get_weights() order and 'th' layouts, so Keras-trained arrays drop in with set_weights().
    """

    def __init__(self, board, input_dim, specs, head="policy", dense_units=256, dtype="fp16",
                 kernel_flip=True, seed=None, unit_end=None):
        """unit_end (residual networks only, policy.py:141-271): one flag per convolution, 1 where it
        closes a residual unit (the first convolution counts as closing one: its output is the
        first skip path).  None = a plain Sequential stack."""
        self.board, self.input_dim = int(board), int(input_dim)
        self.specs = [(int(f), int(w)) for f, w in specs]
        self.unit_end = None if unit_end is None else [int(bool(u)) for u in unit_end]
        if self.unit_end is not None and (len(self.unit_end) != len(self.specs) or not self.unit_end[0] or not self.unit_end[-1]):
            raise ValueError("unit_end needs one flag per convolution, set on the first and the last")
        self.head, self.dense_units = head, int(dense_units)
        if dtype not in DTYPE_CODES:
            raise ValueError("dtype must be one of %s (got %r)" % (sorted(DTYPE_CODES), dtype))
        self.dtype, self.kernel_flip = dtype, bool(kernel_flip)
        self.input_planes = None        # canonical plane of every input plane (set by NeuralNetBase): a ranking hint for 'fp16c32/64'
        self.input_shape = (None, self.input_dim, self.board, self.board)      # read by trainers
        self.output_shape = (None, self.board * self.board) if head == "policy" else (None, 1)
        self._weights = self._init_uniform(seed)
        self._handle = None
        self._handle_device = None
        self._dirty = True
        self._ws = None
        self.uses_learning_phase = False

    # -- weights ---------------------------------------------------------------------------
    def _shapes(self):
        shapes, cin = [], self.input_dim
        for i, (nb, k) in enumerate(self.specs):
            if self.unit_end is not None and i > 0:
                shapes += [(self.board,)] * 4       # BatchNormalization: gamma, beta, running_mean, running_std (axis -1)
            shapes += [(nb, cin, k, k), (nb,)]
            cin = nb
        shapes += [(1, cin, 1, 1), (1,)]
        P = self.board * self.board
        if self.head == "policy":
            shapes += [(P,)]
        else:
            shapes += [(P, self.dense_units), (self.dense_units,), (self.dense_units, 1), (1,)]
        return shapes

    def _init_uniform(self, seed):
        """init='uniform' (Keras 1.2.0: U(-0.05, 0.05)); biases and Bias start at zero."""
        rng = np.random.RandomState(seed)
        out = [rng.uniform(-0.05, 0.05, size=s).astype(np.float32) if len(s) > 1 else np.zeros(s, np.float32)
               for s in self._shapes()]
        if self.unit_end is not None:               # BatchNormalization starts as gamma = 1, beta = 0, mean 0, variance 1
            i = 2
            for _ in self.specs[1:]:
                out[i] = np.ones(self.board, np.float32)
                out[i + 3] = np.ones(self.board, np.float32)
                i += 6
        return out

    def get_weights(self):
        return [w.copy() for w in self._weights]

    def set_weights(self, weights):
        shapes = self._shapes()
        if len(weights) != len(shapes):
            raise ValueError("expected %d weight arrays, got %d" % (len(shapes), len(weights)))
        new = []
        for w, s in zip(weights, shapes):
            w = np.ascontiguousarray(w, dtype=np.float32)
            if tuple(w.shape) != tuple(s):
                raise ValueError("weight shape %s does not match %s" % (w.shape, s))
            new.append(w)
        self._weights = new
        self._dirty = True

    def _keras_layers(self):
        """[(layer name, [weight names])] in the order Keras 1.x stores them (`layer_names` /
        `weight_names` attributes of a weight file); the weights follow `get_weights()` order."""
        layers = []
        for i in range(len(self.specs)):
            if self.unit_end is not None and i > 0:
                bn = "batchnormalization_%d" % i
                layers.append((bn, [bn + "_gamma", bn + "_beta", bn + "_running_mean", bn + "_running_std"]))
            cv = "convolution2d_%d" % (i + 1)
            layers.append((cv, [cv + "_W", cv + "_b"]))
        cv = "convolution2d_%d" % (len(self.specs) + 1)
        layers += [(cv, [cv + "_W", cv + "_b"]), ("flatten_1", [])]
        if self.head == "policy":
            layers += [("bias_1", ["param_0"]), ("activation_1", [])]
        else:
            layers += [("dense_1", ["dense_1_W", "dense_1_b"]), ("dense_2", ["dense_2_W", "dense_2_b"])]
        return layers

    def save_weights(self, path, overwrite=True):
        """Keras 1.x weight file (HDF5: root attribute `layer_names`, one group per layer with a
        `weight_names` attribute and one dataset per array) when `path` ends in .hdf5/.h5 -- what
        nn_util.py:93-96 writes through `model.save_weights`; any other extension gives an .npz
        with the same arrays in `get_weights()` order."""
        if not overwrite and os.path.exists(path):
            raise IOError("%s exists" % path)
        if not path.lower().endswith((".hdf5", ".h5")):
            with open(path, "wb") as f:
                np.savez(f, **{"w%03d" % i: w for i, w in enumerate(self._weights)})
            return
        w = hdf5_lite.Writer()
        layers = self._keras_layers()
        w.attrs["layer_names"] = np.array([n.encode() for n, _ in layers])
        it = iter(self._weights)
        for name, wnames in layers:
            g = w.create_group(name)
            g.attrs["weight_names"] = np.array([n.encode() for n in wnames]) if wnames else np.zeros(0, "S1")
            for wn in wnames:
                g[wn] = next(it)
        w.save(path)

    def load_weights(self, path):
        """Reads a Keras 1.x HDF5 weight file (e.g. the reference's
        tests/test_data/hdf5/random_minimodel_weights.hdf5) or this class's .npz."""
        with open(path, "rb") as f:
            magic = f.read(8)
        if magic != hdf5_lite.SIGNATURE:
            z = np.load(path)
            self.set_weights([z["w%03d" % i] for i in range(len(z.files))])
            return
        f = hdf5_lite.File(path)
        if "layer_names" not in f.attrs and "model_weights" in f:
            f = f["model_weights"]                          # a full `model.save()` file
        if "layer_names" not in f.attrs:
            raise ValueError("%s has no layer_names attribute: not a Keras weight file" % path)
        arrays = []
        for layer in f.attrs["layer_names"]:
            g = f[layer.decode("utf-8")]
            for wn in np.atleast_1d(g.attrs.get("weight_names", [])):
                if isinstance(wn, bytes):
                    arrays.append(g[wn.decode("utf-8")][...])
        self.set_weights(arrays)

    # -- (de)serialisation in Keras-1 Sequential JSON -------------------------------------------
    def _functional_json(self):
        """Keras 1.2.0 `Model.to_json()` of ResnetPolicy.create_network's graph (policy.py:196-271): an
        InputLayer, then layers with `inbound_nodes`, Merge(mode='sum') closing every residual unit."""
        layers, counter = [], {}

        def add(cls, cfg, inbound):
            counter[cls] = counter.get(cls, 0) + 1
            name = "%s_%d" % ({"InputLayer": "input", "Convolution2D": "convolution2d", "BatchNormalization": "batchnormalization",
                               "Activation": "activation", "Merge": "merge", "Flatten": "flatten", "Bias": "bias"}[cls], counter[cls])
            cfg = dict(cfg, name=name)
            layers.append({"class_name": cls, "config": cfg, "name": name,
                           "inbound_nodes": [[[src, 0, 0] for src in inbound]] if inbound else []})
            return name

        def conv(nb, k, inbound):
            return add("Convolution2D", {"nb_filter": nb, "nb_row": k, "nb_col": k, "init": "uniform", "activation": "linear",
                                         "border_mode": "same", "dim_ordering": "th", "subsample": [1, 1], "bias": True,
                                         "trainable": True}, [inbound])
        inp = add("InputLayer", {"batch_input_shape": [None, self.input_dim, self.board, self.board], "input_dtype": "float32",
                                 "sparse": False}, [])
        path = conv(self.specs[0][0], self.specs[0][1], inp)
        block_input = path
        for i in range(1, len(self.specs)):
            path = add("BatchNormalization", {"epsilon": 0.001, "mode": 0, "axis": -1, "momentum": 0.99, "trainable": True}, [path])
            path = add("Activation", {"activation": "relu", "trainable": True}, [path])
            path = conv(self.specs[i][0], self.specs[i][1], path)
            if self.unit_end[i]:
                counter["Merge"] = counter.get("Merge", 0) + 1
                name = "merge_%d" % counter["Merge"]
                layers.append({"class_name": "Merge", "name": name, "inbound_nodes": [[[block_input, 0, 0], [path, 0, 0]]],
                               "config": {"name": name, "mode": "sum", "concat_axis": -1, "dot_axes": -1, "mode_type": "raw",
                                          "output_shape": None, "output_shape_type": "raw", "output_mask": None,
                                          "output_mask_type": "raw", "arguments": {}}})
                path = block_input = name
        path = add("Activation", {"activation": "relu", "trainable": True}, [path])
        path = add("Convolution2D", {"nb_filter": 1, "nb_row": 1, "nb_col": 1, "init": "uniform", "activation": "linear",
                                     "border_mode": "same", "dim_ordering": "th", "subsample": [1, 1], "bias": True,
                                     "trainable": True}, [path])
        path = add("Flatten", {"trainable": True}, [path])
        path = add("Bias", {"trainable": True}, [path])
        path = add("Activation", {"activation": "softmax", "trainable": True}, [path])
        return {"class_name": "Model", "keras_version": "1.2.0",
                "config": {"name": "model_1", "layers": layers, "input_layers": [[inp, 0, 0]], "output_layers": [[path, 0, 0]]},
                "rocalphago_b200": {"dtype": self.dtype, "kernel_flip": self.kernel_flip}}

    @staticmethod
    def _from_functional(km):
        """Parses the functional-model JSON Keras 1.x writes for a reference-built ResnetPolicy
        (policy.py:271 `Model(input=..., output=...)`, saved by nn_util.py:96-101): layers come in
        topological order; a Merge(mode='sum') after a convolution marks the end of a residual unit."""
        cfg = km["config"]
        by_name = {l["name"]: l for l in cfg["layers"]}
        convs, unit_end, board, input_dim, saw_softmax = [], [], None, None, False
        pending_bn = False
        for layer in cfg["layers"]:
            cls, c = layer["class_name"], layer["config"]
            if cls == "InputLayer":
                _, input_dim, board, _ = c["batch_input_shape"]
            elif cls == "Convolution2D":
                if c.get("border_mode", "same") != "same" or tuple(c.get("subsample", (1, 1))) != (1, 1) or c["nb_row"] != c["nb_col"]:
                    raise ValueError("only square stride-1 'same' convolutions are supported")
                if c.get("activation", "linear") != "linear":
                    raise ValueError("convolutions of a residual network are linear (policy.py:207,239)")
                if convs and not (c["nb_filter"] == 1 and c["nb_row"] == 1) and not pending_bn:
                    raise ValueError("expected BatchNormalization -> ReLU in front of %s" % layer["name"])
                convs.append((c["nb_filter"], c["nb_row"]))
                unit_end.append(1 if len(convs) == 1 else 0)
                pending_bn = False
            elif cls == "BatchNormalization":
                if c.get("mode", 0) != 0 or c.get("axis", -1) not in (-1, 3):
                    raise ValueError("BatchNormalization(mode=%r, axis=%r) is not what ResnetPolicy builds" % (c.get("mode"), c.get("axis")))
                pending_bn = True
            elif cls == "Merge":
                srcs = [s[0] for s in layer["inbound_nodes"][0]]
                if c.get("mode") != "sum" or len(srcs) != 2 or not all(s in by_name for s in srcs):
                    raise ValueError("only Merge(mode='sum') of a unit's input and its path is supported")
                unit_end[-1] = 1
            elif cls == "Activation":
                if c.get("activation") == "softmax":
                    saw_softmax = True
                elif c.get("activation") != "relu":
                    raise ValueError("activation %r is not supported on this path" % c.get("activation"))
            elif cls in ("Flatten", "Bias"):
                pass
            else:
                raise ValueError("layer %s is not supported on this path" % cls)
        if len(convs) < 2 or convs[-1] != (1, 1) or not saw_softmax:
            raise ValueError("expected the graph to end in a 1x1 convolution, Flatten, Bias and softmax")
        ext = km.get("rocalphago_b200", {})
        return Network(board, input_dim, convs[:-1], head="policy", dtype=ext.get("dtype", "fp16"),
                       kernel_flip=ext.get("kernel_flip", True), unit_end=unit_end[:-1])

    def to_json(self):
        if self.unit_end is not None:
            return json.dumps(self._functional_json())
        layers = []
        for i, (nb, k) in enumerate(self.specs):
            cfg = {"name": "convolution2d_%d" % (i + 1), "nb_filter": nb, "nb_row": k, "nb_col": k,
                   "init": "uniform", "activation": "relu", "border_mode": "same", "dim_ordering": "th",
                   "subsample": [1, 1], "bias": True}
            if i == 0:
                cfg["batch_input_shape"] = [None, self.input_dim, self.board, self.board]
            layers.append({"class_name": "Convolution2D", "config": cfg})
        layers.append({"class_name": "Convolution2D", "config": {
            "name": "convolution2d_%d" % (len(self.specs) + 1), "nb_filter": 1, "nb_row": 1, "nb_col": 1,
            "init": "uniform", "activation": "linear", "border_mode": "same", "dim_ordering": "th",
            "subsample": [1, 1], "bias": True}})
        layers.append({"class_name": "Flatten", "config": {"name": "flatten_1"}})
        if self.head == "policy":
            layers.append({"class_name": "Bias", "config": {"name": "bias_1"}})
            layers.append({"class_name": "Activation", "config": {"activation": "softmax", "name": "activation_1"}})
        else:
            layers.append({"class_name": "Dense", "config": {"name": "dense_1", "output_dim": self.dense_units,
                                                             "init": "uniform", "activation": "linear"}})
            layers.append({"class_name": "Dense", "config": {"name": "dense_2", "output_dim": 1,
                                                             "init": "uniform", "activation": "tanh"}})
        return json.dumps({"class_name": "Sequential", "keras_version": "1.2.0", "config": layers,
                           "rocalphago_b200": {"dtype": self.dtype, "kernel_flip": self.kernel_flip}})

    @staticmethod
    def from_json(text):
        """Reads the Sequential JSON Keras 1.x writes for these architectures
        (e.g. the reference's tests/test_data/minimodel.json)."""
        km = json.loads(text)
        ext = km.get("rocalphago_b200", {})
        if ext.get("arch") == "resnet":
            return Network(ext["board"], ext["input_dim"], [tuple(x) for x in ext["specs"]], head="policy",
                           dtype=ext.get("dtype", "fp16"), kernel_flip=ext.get("kernel_flip", True),
                           unit_end=ext["unit_end"])
        if km.get("class_name") == "Model":
            return Network._from_functional(km)
        if km.get("class_name") != "Sequential":
            raise ValueError("only Sequential and functional (Model) JSON is supported on this path (got %s)" % km.get("class_name"))
        convs, head, dense, board, input_dim = [], None, [], None, None
        for layer in km["config"]:
            name, cfg = layer["class_name"], layer["config"]
            if name == "Convolution2D":
                if cfg.get("border_mode", "same") != "same" or tuple(cfg.get("subsample", (1, 1))) != (1, 1):
                    raise ValueError("only stride-1 'same' convolutions are supported")
                if cfg["nb_row"] != cfg["nb_col"]:
                    raise ValueError("only square filters are supported")
                if "batch_input_shape" in cfg:
                    _, input_dim, board, _ = cfg["batch_input_shape"]
                convs.append((cfg["nb_filter"], cfg["nb_row"], cfg.get("activation", "linear")))
            elif name == "Dense":
                dense.append((cfg["output_dim"], cfg.get("activation", "linear")))
            elif name == "Activation" and cfg.get("activation") == "softmax":
                head = "policy"
            elif name in ("Flatten", "Bias"):
                pass
            else:
                raise ValueError("layer %s is not supported on this path" % name)
        if not convs or convs[-1][:2] != (1, 1) or convs[-1][2] != "linear":
            raise ValueError("expected the stack to end in a linear 1x1 convolution to one plane")
        if any(a != "relu" for _, _, a in convs[:-1]):
            raise ValueError("hidden convolutions must use relu")
        if head is None:
            if len(dense) != 2 or dense[1] != (1, "tanh") or dense[0][1] != "linear":
                raise ValueError("expected Dense(n, linear) + Dense(1, tanh) as the value head")
            head = "value"
        extra = km.get("rocalphago_b200", {})
        return Network(board, input_dim, [(f, w) for f, w, _ in convs[:-1]], head=head,
                       dense_units=dense[0][0] if dense else 256, dtype=extra.get("dtype", "fp16"),
                       kernel_flip=extra.get("kernel_flip", True))

    # -- the CUDA forward ---------------------------------------------------------------------
    def _ensure_handle(self, device):
        L = _cabi.lib()
        if self._handle is not None and self._handle_device != device:
            L.rag_net_destroy(self._handle)
            self._handle = None
        if self._handle is None:
            if len(self.specs) > MAX_LAYERS:
                raise ValueError("at most %d layers are supported" % MAX_LAYERS)
            d = _NetDesc()
            d.board, d.in_planes, d.layers = self.board, self.input_dim, len(self.specs)
            for i, (nb, k) in enumerate(self.specs):
                d.filters[i], d.widths[i] = nb, k
            d.head = HEAD_POLICY if self.head == "policy" else HEAD_VALUE
            d.dense_units = self.dense_units
            d.dtype = DTYPE_CODES[self.dtype]
            d.kernel_flip = int(self.kernel_flip)
            d.resnet = int(self.unit_end is not None)
            for i, u in enumerate(self.unit_end or []):
                d.unit_end[i] = u
            h = ctypes.c_void_p()
            with torch.cuda.device(device):
                _cabi.check(L.rag_net_create(ctypes.byref(d), ctypes.byref(h)))
            self._handle, self._handle_device, self._dirty = h, device, True
            if self.input_planes is not None and len(self.input_planes) == self.input_dim:
                src = np.ascontiguousarray(self.input_planes, np.int32)
                _cabi.check(L.rag_net_set_input_planes(h, src.ctypes.data_as(ctypes.c_void_p), len(src)))
        if self._dirty:
            arr = (ctypes.c_void_p * len(self._weights))(*[w.ctypes.data_as(ctypes.c_void_p).value for w in self._weights])
            with torch.cuda.device(device):
                _cabi.check(L.rag_net_load_weights(self._handle, arr, len(self._weights)))
            self._dirty = False
        return self._handle

    def _workspace(self, nbytes, device):
        if self._ws is None or self._ws.numel() < nbytes or self._ws.device != device:
            self._ws = torch.empty(int(nbytes), dtype=torch.uint8, device=device)
        return self._ws

    def _out(self, n, device):
        return torch.empty((n, self.board * self.board) if self.head == "policy" else (n,),
                           dtype=torch.float32, device=device)

    def forward_device(self, x):
        """x: CUDA float32 (N, F, S, S) -> CUDA float32 (N, S*S) or (N,)."""
        _require_cuda()
        n = int(x.shape[0])
        if tuple(x.shape[1:]) != (self.input_dim, self.board, self.board):
            raise ValueError("input shape %s does not match %s" % (tuple(x.shape), self.input_shape))
        x = x.contiguous()
        h = self._ensure_handle(x.device)
        out = self._out(n, x.device)
        if n == 0:
            return out
        L = _cabi.lib()
        ws = self._workspace(L.rag_net_workspace_bytes(h, n), x.device)
        with torch.cuda.device(x.device):
            _cabi.check(L.rag_net_forward(h, _ptr(x), n, _ptr(out), _ptr(ws), ws.numel(), _stream()))
        return out

    def eval_batch_device(self, batch, plane_src):
        """Packed states -> features -> forward without leaving the device (rag_eval_batch)."""
        _require_cuda()
        n = len(batch)
        h = self._ensure_handle(batch.device)
        out = self._out(n, batch.device)
        if n == 0:
            return out
        L = _cabi.lib()
        src = np.ascontiguousarray(plane_src, np.int32)
        if getattr(batch, "superko", None) is not None:
            # positional superko needs the analysis entry point that knows the move history; the rest
            # is what rag_eval_batch does: conv-layout emit, then the forward over pixel rows
            vp = ctypes.c_void_p
            key = (n, batch.device)
            if getattr(self, "_sk_key", None) != key:            # the three buffers of this branch are kept per batch size
                self._sk_bufs = (torch.empty(int(L.rag_features_workspace_bytes(n)), dtype=torch.uint8, device=batch.device),
                                 torch.empty((int(L.rag_net_input_rows(h, n)), int(L.rag_net_input_channels(h))), dtype=torch.float16,
                                             device=batch.device),
                                 torch.empty(int(L.rag_net_workspace_bytes(h, n)), dtype=torch.uint8, device=batch.device))
                self._sk_key = key
            fws, rows, nws = self._sk_bufs
            self._ws = fws
            analyze_into(batch, fws, True)
            with torch.cuda.device(batch.device):
                _cabi.check(L.rag_features_emit(_ptr(fws), n, batch.size, src.ctypes.data_as(vp), len(src), LAYOUT_CONV_IN,
                                                DTYPE_BF16 if self.dtype == "bf16" else DTYPE_FP16, rows.shape[1],
                                                _ptr(rows), _stream()))
                _cabi.check(L.rag_net_forward_rows(h, _ptr(rows), n, _ptr(out), _ptr(nws), nws.numel(), _stream()))
            return out
        ws = self._workspace(L.rag_eval_workspace_bytes(h, n), batch.device)
        with torch.cuda.device(batch.device):
            _cabi.check(L.rag_eval_batch(h, _ptr(batch.packed), n, src.ctypes.data_as(ctypes.c_void_p), len(src),
                                         _ptr(out), _ptr(ws), ws.numel(), _stream()))
        return out

    def last_logits(self, n):
        """fp32 output of the 1x1 convolution of the last forward_device(n boards) call."""
        off = int(_cabi.lib().rag_net_logits_offset(self._handle, n))
        P = self.board * self.board
        return self._ws[off:off + 4 * n * P].view(torch.float32).reshape(n, P).clone()

    def features_workspace(self):
        """The analysis workspace of the last eval_batch_device call (planes 46/48 give the move lists)."""
        return self._ws

    def __del__(self):
        try:
            if self._handle is not None:
                _cabi.lib().rag_net_destroy(self._handle)
        except Exception:  # noqa: BLE001
            pass


class NeuralNetBase(object):
    """Base class for neural network classes handling feature processing, construction
    of a 'forward' function, etc. (nn_util.py:8-108)"""

    subclasses = {}

    def __init__(self, feature_list, **kwargs):
        self.preprocessor = Preprocess(feature_list)
        kwargs["input_dim"] = self.preprocessor.output_dim
        if kwargs.get("init_network", True):
            self.model = self.__class__.create_network(**kwargs)
            self.model.input_planes = list(self.preprocessor.plane_src)
            self.forward = self._model_forward()

    def _model_forward(self):
        """forward(ndarray (N,F,S,S)) -> ndarray float32, as nn_util.py:34-58's K.function closure."""
        def forward(inpt):
            _require_cuda()
            x = torch.as_tensor(np.ascontiguousarray(inpt, dtype=np.float32)).cuda(non_blocking=True)
            return self.model.forward_device(x).cpu().numpy().reshape((x.shape[0],) + self.model.output_shape[1:])
        return forward

    @staticmethod
    def load_model(json_file):
        with open(json_file, "r") as f:
            object_specs = json.load(f)
        class_name = object_specs.get("class", "CNNPolicy")
        try:
            network_class = NeuralNetBase.subclasses[class_name]
        except KeyError:
            raise ValueError("Unknown neural network type in json file: {}\n"
                             "(was it registered with the @neuralnet decorator?)".format(class_name))
        new_net = network_class(object_specs["feature_list"], init_network=False)
        new_net.model = Network.from_json(object_specs["keras_model"])
        new_net.model.input_planes = list(new_net.preprocessor.plane_src)
        if "weights_file" in object_specs:
            new_net.model.load_weights(_find_relative(object_specs["weights_file"], json_file))
        new_net.forward = new_net._model_forward()
        return new_net

    def save_model(self, json_file, weights_file=None):
        object_specs = {
            "class": self.__class__.__name__,
            "keras_model": self.model.to_json(),
            "feature_list": self.preprocessor.feature_list,
        }
        if weights_file is not None:
            self.model.save_weights(weights_file)
            object_specs["weights_file"] = weights_file
        with open(json_file, "w") as f:
            json.dump(object_specs, f)


def _find_relative(path, json_file):
    """The reference resolves `weights_file` against the working directory (its files say e.g.
    'tests/test_data/hdf5/...' and are loaded from the repository root).  Same here; if that fails,
    the directories above the JSON file are tried, so that a model can be loaded from anywhere."""
    if os.path.isabs(path) or os.path.exists(path):
        return path
    d = os.path.dirname(os.path.abspath(json_file))
    while True:
        cand = os.path.join(d, path)
        if os.path.exists(cand):
            return cand
        parent = os.path.dirname(d)
        if parent == d:
            return path
        d = parent


def neuralnet(cls):
    """Class decorator for registering subclasses of NeuralNetBase (nn_util.py:111-115)"""
    NeuralNetBase.subclasses[cls.__name__] = cls
    return cls
