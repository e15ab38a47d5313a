"""`clip`-compatible face of the B200 encoder (the subset of openai/CLIP's API the reference calls).

The reference does `import clip` (un-vendored submodule Backend/CLIP = openai/CLIP @ dcba3cb) and uses:
  clip.load(name, device=..., jit=False) -> (model, preprocess)   embedding.py:22, embedding_service.py:86,106
  model.encode_image(Tensor[B,3,224,224]) -> Tensor[B,D]           embedding.py:49, embedding_service.py:41,414,490
  model.visual.output_dim, model.logit_scale, model.dtype, .float(), .eval(), .to()   embedding_service.py:22-34
  clip.tokenize / model.encode_text                                 embedding_service.py:169-177 (text.TextEncoder)
`load` here returns a model whose `encode_image` runs the hand-written sm_100a tower (vit.VisionEncoder ->
libevr.so).  Put this module on sys.path as `clip` and the reference's service code runs unchanged.

Weights: openai checkpoints are fetched from the network by the original `clip.load`; this box has none, so `load`
takes them from (1) an explicit `state_dict=`, (2) `<download_root>/<name with / -> ->.pt` -- OpenAI's own file (a
TorchScript archive, opened with `torch.jit.load` as the original does), a torch-saved openai-keyed state_dict, or a
checkpoint dict with "state_dict"/"model_state_dict" -- else (3) raises, unless `random_init=<seed>` (or
EVR_CLIP_RANDOM_INIT=<seed>) asks for seeded openai-scale random weights (benchmarks; announced loudly on stderr).
`clip.load(<path to a checkpoint file>)` reads the geometry off the tensors (`infer_model_name`).
"""
import os

import torch

from .text import TextEncoder, text_weight_keys
from .vit import GEOMETRY, VisionEncoder, weight_keys

CLIP_MEAN = (0.48145466, 0.4578275, 0.40821073)  # clip/clip.py _transform; compare_models.py:39
CLIP_STD = (0.26862954, 0.26130258, 0.27577711)

_MODELS = tuple(GEOMETRY)


def available_models():
    return list(_MODELS)


def random_visual_state_dict(name="ViT-B/32", seed=1):
    """Seeded random vision-tower weights with openai's init scales (clip/model.py VisionTransformer.__init__ and
    CLIP.initialize_parameters): conv1/class/pos/proj ~ N(0, d^-0.5), in_proj std d^-0.5, out_proj std
    d^-0.5 (2 layers)^-0.5, c_fc std (2d)^-0.5, LayerNorm gamma 1 beta 0.  Keys carry the "visual." prefix."""
    g = GEOMETRY[name]
    d, layers, p, out = g["width"], g["layers"], g["patch"], g["out_dim"]
    L = (g["image_size"] // p) ** 2 + 1
    gen = torch.Generator().manual_seed(int(seed))

    def n(*shape, std):
        return torch.randn(*shape, generator=gen) * std

    s = d ** -0.5
    sd = {"conv1.weight": n(d, 3, p, p, std=s), "class_embedding": n(d, std=s),
          "positional_embedding": n(L, d, std=s), "ln_pre.weight": torch.ones(d), "ln_pre.bias": torch.zeros(d)}
    for i in range(layers):
        b = f"transformer.resblocks.{i}."
        sd[b + "ln_1.weight"], sd[b + "ln_1.bias"] = torch.ones(d), torch.zeros(d)
        sd[b + "attn.in_proj_weight"] = n(3 * d, d, std=s)
        sd[b + "attn.in_proj_bias"] = torch.zeros(3 * d)
        sd[b + "attn.out_proj.weight"] = n(d, d, std=s * (2 * layers) ** -0.5)
        sd[b + "attn.out_proj.bias"] = torch.zeros(d)
        sd[b + "ln_2.weight"], sd[b + "ln_2.bias"] = torch.ones(d), torch.zeros(d)
        sd[b + "mlp.c_fc.weight"] = n(4 * d, d, std=(2 * d) ** -0.5)
        sd[b + "mlp.c_fc.bias"] = torch.zeros(4 * d)
        sd[b + "mlp.c_proj.weight"] = n(d, 4 * d, std=s * (2 * layers) ** -0.5)
        sd[b + "mlp.c_proj.bias"] = torch.zeros(d)
    sd["ln_post.weight"], sd["ln_post.bias"] = torch.ones(d), torch.zeros(d)
    sd["proj"] = n(d, out, std=s)
    full = {"visual." + k: v for k, v in sd.items()}
    full["logit_scale"] = torch.tensor(2.6592)  # ln(1/0.07), clip/model.py
    return full


TEXT_GEOMETRY = {  # clip/model.py build_model: text half
    "ViT-B/32": dict(width=512, layers=12, out_dim=512), "ViT-B/16": dict(width=512, layers=12, out_dim=512),
    "ViT-L/14": dict(width=768, layers=12, out_dim=768),
}


def random_text_state_dict(name="ViT-B/32", seed=1, context_length=77, vocab_size=49408):
    """Seeded random text-tower weights with openai's init scales (CLIP.initialize_parameters: token 0.02,
    positional 0.01, attention d^-0.5, out/c_proj d^-0.5 (2 layers)^-0.5, c_fc (2d)^-0.5, text_projection d^-0.5)."""
    g = TEXT_GEOMETRY[name]
    d, layers, out = g["width"], g["layers"], g["out_dim"]
    gen = torch.Generator().manual_seed(int(seed) + 7919)

    def n(*shape, std):
        return torch.randn(*shape, generator=gen) * std

    s = d ** -0.5
    sd = {"token_embedding.weight": n(vocab_size, d, std=0.02), "positional_embedding": n(context_length, d, std=0.01)}
    for i in range(layers):
        b = f"transformer.resblocks.{i}."
        sd[b + "ln_1.weight"], sd[b + "ln_1.bias"] = torch.ones(d), torch.zeros(d)
        sd[b + "attn.in_proj_weight"], sd[b + "attn.in_proj_bias"] = n(3 * d, d, std=s), torch.zeros(3 * d)
        sd[b + "attn.out_proj.weight"] = n(d, d, std=s * (2 * layers) ** -0.5)
        sd[b + "attn.out_proj.bias"] = torch.zeros(d)
        sd[b + "ln_2.weight"], sd[b + "ln_2.bias"] = torch.ones(d), torch.zeros(d)
        sd[b + "mlp.c_fc.weight"], sd[b + "mlp.c_fc.bias"] = n(4 * d, d, std=(2 * d) ** -0.5), torch.zeros(4 * d)
        sd[b + "mlp.c_proj.weight"] = n(d, 4 * d, std=s * (2 * layers) ** -0.5)
        sd[b + "mlp.c_proj.bias"] = torch.zeros(d)
    sd["ln_final.weight"], sd["ln_final.bias"] = torch.ones(d), torch.zeros(d)
    sd["text_projection"] = n(d, out, std=s)
    return sd


def _find_checkpoint(name, download_root):
    root = download_root or os.environ.get("EVR_CLIP_ROOT") or os.path.expanduser("~/.cache/clip")
    path = os.path.join(root, name.replace("/", "-") + ".pt")
    return path if os.path.isfile(path) else None


def _load_checkpoint_file(path):
    """A weights file -> whatever it holds (module or dict), the way openai's `clip.load` opens one: OpenAI's
    `ViT-B-32.pt` / `ViT-L-14.pt` are TorchScript archives (`torch.jit.load(...).state_dict()`), fine-tuned checkpoints
    (`torch.save({'model_state_dict': ...})`, embedding_service.py:112-113) are pickled dicts.  Pickles are opened with
    `weights_only=True`; a checkpoint that needs arbitrary unpickling is refused unless EVR_CLIP_TRUSTED_CHECKPOINTS=1."""
    import zipfile
    is_script = False
    if zipfile.is_zipfile(path):
        with zipfile.ZipFile(path) as z:
            is_script = any(n.endswith("constants.pkl") for n in z.namelist())
    if is_script:
        return torch.jit.load(path, map_location="cpu").eval().state_dict()
    try:
        return torch.load(path, map_location="cpu", weights_only=True)
    except Exception as e:
        if os.environ.get("EVR_CLIP_TRUSTED_CHECKPOINTS") == "1":
            return torch.load(path, map_location="cpu", weights_only=False)
        raise RuntimeError(f"{path}: not loadable as a TorchScript archive or a plain tensor checkpoint ({e}); set "
                           "EVR_CLIP_TRUSTED_CHECKPOINTS=1 to unpickle a trusted file") from e


def infer_model_name(state_dict):
    """The geometry a state_dict belongs to, read off its tensors the way openai's `build_model` does (vision width
    = conv1 rows, patch = conv1 kernel, layers = resblocks with an in_proj_weight): -> a key of GEOMETRY."""
    conv = state_dict.get("visual.conv1.weight")
    if conv is None:
        raise RuntimeError("checkpoint holds no vision tower (visual.conv1.weight)")
    width, patch = int(conv.shape[0]), int(conv.shape[-1])
    layers = len({k.split(".")[3] for k in state_dict
                  if k.startswith("visual.transformer.resblocks.") and k.endswith(".attn.in_proj_weight")})
    for name, g in GEOMETRY.items():
        if (g["width"], g["patch"], g["layers"]) == (width, patch, layers):
            return name
    raise RuntimeError(f"unsupported vision tower: width {width}, patch {patch}, {layers} layers; available models = "
                       f"{available_models()}")


def _unwrap(obj):
    if hasattr(obj, "state_dict") and not isinstance(obj, dict):
        obj = obj.state_dict()
    for key in ("state_dict", "model_state_dict"):
        if isinstance(obj, dict) and key in obj and isinstance(obj[key], dict):
            obj = obj[key]
    # fine-tuned checkpoints wrap CLIP as `clip_model.` (embedding_service.py:16-24, :113)
    if any(k.startswith("clip_model.") for k in obj):
        obj = {k[len("clip_model."):]: v for k, v in obj.items() if k.startswith("clip_model.")}
    return obj


class _Visual:
    """Stand-in for `model.visual`: the attributes the reference reads (embedding_service.py:25)."""

    def __init__(self, encoder, name):
        g = GEOMETRY[name]
        self.encoder = encoder
        self.output_dim = g["out_dim"]
        self.input_resolution = g["image_size"]

    def __call__(self, image):
        return self.encoder.encode(image, normalise=False)


class CLIPModel:
    """What `clip.load` returns: encode_image on the sm_100a tower; openai's dtype rule (fp16 on CUDA until
    `.float()`, clip.py `convert_weights` / embedding_service.py:22)."""

    def __init__(self, name, state_dict, device, max_batch=256):
        self.name = name
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("this CLIP runs on a CUDA device only (no CPU path); got device=%r" % (device,))
        self._sd = {k: v.detach().to("cpu", torch.float32) for k, v in state_dict.items()
                    if torch.is_tensor(v) and v.is_floating_point()}
        self._text = None
        self._max_batch = max_batch
        enc = VisionEncoder.from_state_dict(self._sd, name, self.device, max_batch=max_batch, prefix="visual.")
        self.visual = _Visual(enc, name)
        self.logit_scale = self._sd.get("logit_scale", torch.tensor(2.6592)).to(self.device)
        self.dtype = torch.float16
        self.training = False

    # ---- nn.Module-like surface the reference touches --------------------------------------------
    def eval(self):
        return self

    def train(self, mode=True):
        if mode:
            raise RuntimeError("the sm_100a tower is inference-only (the reference serves under torch.no_grad())")
        return self

    def float(self):
        self.dtype = torch.float32
        return self

    def half(self):
        self.dtype = torch.float16
        return self

    def to(self, *args, **kwargs):
        for a in list(args) + list(kwargs.values()):
            if isinstance(a, (str, torch.device)) and torch.device(a) != self.device and \
                    not (torch.device(a).type == "cuda" and torch.device(a).index is None):
                raise RuntimeError(f"model lives on {self.device}; reload with clip.load(..., device={a!r})")
            if a in (torch.float32, torch.float16):
                self.dtype = a
        return self

    def parameters(self):
        return iter(self._sd.values())

    def state_dict(self):
        return dict(self._sd)

    @torch.no_grad()
    def encode_image(self, image):
        """[B,3,S,S] float (clip's preprocess output) or [B,S,S,3] uint8 RGB -> [B,D] in the model dtype."""
        out = []
        mb = self.visual.encoder.max_batch
        for i in range(0, image.shape[0], mb):
            out.append(self.visual.encoder.encode(image[i:i + mb], normalise=False, out_dtype=self.dtype))
        if not out:
            return torch.empty((0, self.visual.output_dim), dtype=self.dtype, device=self.device)
        return out[0] if len(out) == 1 else torch.cat(out, 0)

    @property
    def has_text_tower(self):
        return "token_embedding.weight" in self._sd and "text_projection" in self._sd

    @torch.no_grad()
    def encode_text(self, text):
        """LongTensor[B, 77] (clip.tokenize) -> [B, D] in the model dtype, on the sm_100a text tower."""
        if not self.has_text_tower:
            raise NotImplementedError("this checkpoint holds no text tower (token_embedding / text_projection keys)")
        if self._text is None:
            self._text = TextEncoder.from_state_dict(self._sd, self.device, max_batch=min(self._max_batch, 64))
        return self._text.encode(text, normalise=False, out_dtype=self.dtype)

    def __call__(self, image, text=None):
        if text is not None:
            return self.encode_text(text)
        return self.encode_image(image)


def _transform(n_px):
    """clip/clip.py `_transform`: Resize(bicubic) -> CenterCrop -> RGB -> ToTensor -> Normalize (host side, PIL)."""
    from torchvision import transforms as T
    t = T.Compose([
        T.Resize(n_px, interpolation=T.InterpolationMode.BICUBIC), T.CenterCrop(n_px),
        lambda im: im.convert("RGB"), T.ToTensor(), T.Normalize(CLIP_MEAN, CLIP_STD)])
    t.evr_n_px = n_px  # lets the ingest loop swap in the bit-identical device path (preprocess.GpuPreprocess)
    return t


def load(name, device="cuda", jit=False, download_root=None, state_dict=None, random_init=None, max_batch=256):
    """clip.load(name, device, jit=False) -> (model, preprocess).  See the module docstring for weight lookup."""
    if name not in _MODELS:
        if os.path.isfile(name):
            state_dict = _unwrap(_load_checkpoint_file(name))
            name = infer_model_name(state_dict)
        else:
            raise RuntimeError(f"Model {name} not found; available models = {available_models()}")
    if state_dict is None:
        path = _find_checkpoint(name, download_root)
        if path is not None:
            state_dict = _unwrap(_load_checkpoint_file(path))
    if state_dict is None:
        seed = random_init if random_init is not None else os.environ.get("EVR_CLIP_RANDOM_INIT")
        if seed is None:
            raise RuntimeError(
                f"no weights for {name}: this box has no network, put an openai-keyed state_dict at "
                f"{os.path.join(download_root or '~/.cache/clip', name.replace('/', '-') + '.pt')} "
                "(or pass state_dict=..., or random_init=<seed> for benchmarking)")
        import sys
        print(f"WARNING: clip.load({name!r}): no checkpoint found -- serving seeded RANDOM weights (seed {int(seed)}); "
              "embeddings are meaningless outside benchmarks and parity tests", file=sys.stderr)
        state_dict = random_visual_state_dict(name, int(seed))
        state_dict.update(random_text_state_dict(name, int(seed)))
    else:
        state_dict = _unwrap(state_dict)
    missing = [k for k in weight_keys(GEOMETRY[name]["layers"], "visual.") if k not in state_dict]
    if missing:
        raise RuntimeError(f"checkpoint lacks {len(missing)} vision-tower keys, e.g. {missing[:3]}")
    model = CLIPModel(name, state_dict, device, max_batch=max_batch)
    return model, _transform(GEOMETRY[name]["image_size"])


def tokenize(texts, context_length=77, truncate=False, bpe_path=None):
    """clip.tokenize(list[str]) -> LongTensor[B, 77] (tokenizer.py: openai's byte-level BPE restated)."""
    from .tokenizer import tokenize as _tokenize
    return _tokenize(texts, context_length, truncate, bpe_path)
