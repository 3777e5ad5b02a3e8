"""ControlNet input pre-processors (reference: swarm/controlnet/input_processor.py:17-272).

Dispatch on `controlnet["type"]` exactly as the reference.  canny / tile / pix2pix / shuffle are
self-contained (OpenCV / PIL / numpy).  The learned annotators (mlsd, depth, normalbae, seg,
lineart, openpose, scribble, softedge) come from `controlnet_aux` / transformers hub models in the
reference; they are used when importable and locally cached; soft edges / scribbles and lineart additionally have this
tree's own networks (controlnet/annotators.py: HED from `lllyasviel/Annotators/ControlNetHED.pth`, the lineart generator
from `sk_model.pth`, PiDiNet from `table5_pidinet.pth` in controlnet/pidinet.py), openpose this tree's body-pose detector (controlnet/openpose.py, `body_pose_model.pth`) and mlsd this
tree's M-LSD network (controlnet/mlsd.py, `mlsd_large_512_fp32.pth`) and normalbae this tree's NNET (controlnet/normalbae.py, `scannet.pt`);
otherwise a
classical CPU approximation keeps the job serviceable (documented per type), so no type raises for a missing
dependency.  Unknown types raise, as in the reference."""
from __future__ import annotations

import logging
import threading

import numpy as np
from PIL import Image

_reported = set()
_tls = threading.local()


def _stand_in(kind: str, what: str):
    """The learned annotator of `kind` is not on this worker: say so (in the log once per type, and on the image the
    pre-processor returns, from where the job callback copies it into `pipeline_config`) instead of degrading silently."""
    _tls.stand_in = what
    if kind not in _reported:
        _reported.add(kind)
        logging.warning("controlnet pre-processor %s: learned annotator not available locally, using %s", kind, what)


def preprocess_image(image, controlnet):
    if not controlnet.get("preprocess", False):
        return image
    kind = controlnet.get("type")
    handlers = {
        "canny": lambda im: image_to_canny(im, controlnet),
        "mlsd": image_to_mlsd,
        "depth": image_to_depth,
        "normalbae": image_to_normal,
        "seg": image_to_segmentation,
        "lineart": image_to_lineart,
        "openpose": image_to_openpose,
        "pix2pix": lambda im: im,
        "scribble": lambda im: image_to_softedge(im, scribble=True),
        "softedge": image_to_softedge,
        "shuffle": image_to_shuffle,
        "tile": image_to_tile,
    }
    if kind not in handlers:
        raise Exception("Unknown controlnet type")
    _tls.stand_in = None
    out = handlers[kind](image)
    if getattr(_tls, "stand_in", None) and hasattr(out, "info"):
        out.info["controlnet_annotator"] = f"stand-in for {kind}: {_tls.stand_in}"
    return out


def _gray3(arr) -> Image.Image:
    arr = np.asarray(arr).astype(np.uint8)
    return Image.fromarray(np.stack([arr] * 3, axis=2))


def image_to_tile(image, resolution=1024):
    rgb = image.convert("RGB")
    w, h = rgb.size
    k = float(resolution) / min(h, w)
    h2 = int(round(h * k / 64.0)) * 64
    w2 = int(round(w * k / 64.0)) * 64
    return rgb.resize((w2, h2), resample=Image.Resampling.LANCZOS)


def image_to_canny(image, controlnet):
    import cv2

    edges = cv2.Canny(np.array(image), controlnet.get("low_threshold", 100), controlnet.get("high_threshold", 200))
    return _gray3(edges)


def _aux(name, repo, **kw):
    """controlnet_aux detector when the package and its weights are available, else None."""
    try:
        import controlnet_aux

        return getattr(controlnet_aux, name).from_pretrained(repo, **kw)
    except Exception:
        return None


def _sobel_mag(image):
    import cv2

    g = cv2.cvtColor(np.array(image.convert("RGB")), cv2.COLOR_RGB2GRAY).astype(np.float32)
    g = cv2.GaussianBlur(g, (5, 5), 0)
    gx, gy = cv2.Sobel(g, cv2.CV_32F, 1, 0), cv2.Sobel(g, cv2.CV_32F, 0, 1)
    mag = np.sqrt(gx * gx + gy * gy)
    return gx, gy, mag / max(float(mag.max()), 1e-6)


def image_to_mlsd(image):
    det = _aux("MLSDdetector", "lllyasviel/ControlNet")
    if det is not None:
        return det(image)
    from .mlsd import load_mlsd  # this tree's own M-LSD network, when its checkpoint is present locally

    own = load_mlsd()
    if own is not None:
        return own(image)
    import cv2  # fallback: probabilistic Hough line segments on Canny edges

    _stand_in("mlsd", "Hough line segments on Canny edges")

    arr = np.array(image.convert("RGB"))
    edges = cv2.Canny(arr, 60, 180)
    canvas = np.zeros(edges.shape, np.uint8)
    lines = cv2.HoughLinesP(edges, 1, np.pi / 180, 40, minLineLength=max(8, min(edges.shape) // 16), maxLineGap=4)
    if lines is not None:
        for x1, y1, x2, y2 in lines[:, 0]:
            cv2.line(canvas, (int(x1), int(y1)), (int(x2), int(y2)), 255, 1)
    return _gray3(canvas)


_learned_cache = {}


def _learned(kind: str, loader):
    """A transformers-backed annotator, built once per process; a failed attempt (not cached locally, no network) is
    remembered too, so later jobs go straight to the stand-in instead of retrying the hub."""
    if kind not in _learned_cache:
        try:
            _learned_cache[kind] = loader()
        except Exception as e:
            logging.info("controlnet pre-processor %s: %s", kind, e)
            _learned_cache[kind] = None
    return _learned_cache[kind]


def image_to_depth(image):
    try:
        def load():
            from transformers import pipeline

            return pipeline("depth-estimation")

        estimator = _learned("depth", load)
        if estimator is None:
            raise RuntimeError("no depth estimator")
        return _gray3(np.array(estimator(image)["depth"]))
    except Exception:
        # fallback: defocus heuristic (sharp = near) so the control image has the right modality
        _stand_in("depth", "a defocus heuristic")
        _, _, mag = _sobel_mag(image)
        import cv2

        d = cv2.GaussianBlur(mag, (0, 0), max(image.size) / 40.0)
        return _gray3(255 * d / max(float(d.max()), 1e-6))


def image_to_normal(image):
    det = _aux("NormalBaeDetector", "lllyasviel/Annotators")
    if det is not None:
        return det(image)
    from .normalbae import load_normalbae  # this tree's own NNET, when its checkpoint is present locally

    nnet = load_normalbae()
    if nnet is not None:
        return nnet.normals(image, next(nnet.parameters()).device)
    _stand_in("normalbae", "normals from Sobel gradients")
    gx, gy, _ = _sobel_mag(image)
    n = np.stack([-gx, -gy, np.full_like(gx, 255.0 / 4)], axis=2)
    n /= np.linalg.norm(n, axis=2, keepdims=True) + 1e-6
    return Image.fromarray(((n * 0.5 + 0.5) * 255).astype(np.uint8))


def image_to_segmentation(image):
    try:
        import torch
        from transformers import AutoImageProcessor, UperNetForSemanticSegmentation

        loaded = _learned("seg", lambda: (AutoImageProcessor.from_pretrained("openmmlab/upernet-convnext-small"),
                                          UperNetForSemanticSegmentation.from_pretrained("openmmlab/upernet-convnext-small").eval()))
        if loaded is None:
            raise RuntimeError("no segmentation model")
        proc, seg_model = loaded
        with torch.no_grad():
            out = seg_model(proc(image, return_tensors="pt").pixel_values)
        seg = proc.post_process_semantic_segmentation(out, target_sizes=[image.size[::-1]])[0].numpy()
    except Exception:
        import cv2  # fallback: k-means colour clustering into palette classes

        _stand_in("seg", "k-means colour clustering")

        arr = np.array(image.convert("RGB"))
        small = cv2.resize(arr, (max(8, arr.shape[1] // 4), max(8, arr.shape[0] // 4)))
        z = small.reshape(-1, 3).astype(np.float32)
        k = 8
        _, labels, _ = cv2.kmeans(z, k, None, (cv2.TERM_CRITERIA_EPS + cv2.TERM_CRITERIA_MAX_ITER, 10, 1.0), 2,
                                  cv2.KMEANS_PP_CENTERS)
        seg = cv2.resize(labels.reshape(small.shape[:2]).astype(np.uint8), (arr.shape[1], arr.shape[0]),
                         interpolation=cv2.INTER_NEAREST)
    color = np.zeros((*seg.shape, 3), dtype=np.uint8)
    palette = ade_palette()
    for label in np.unique(seg):
        color[seg == label] = palette[int(label) % len(palette)]
    return Image.fromarray(color)


def image_to_lineart(image):
    det = _aux("LineartDetector", "lllyasviel/Annotators")
    if det is not None:
        return det(image)
    from .annotators import load_lineart  # this tree's own lineart generator, when its checkpoint is present locally

    net = load_lineart()
    if net is not None:
        return net.lines(image, next(net.parameters()).device)
    _stand_in("lineart", "Sobel edge magnitude")
    _, _, mag = _sobel_mag(image)
    return _gray3(255 * np.clip(mag * 3.0, 0, 1))


def image_to_openpose(image):
    det = _aux("OpenposeDetector", "lllyasviel/ControlNet")
    if det is not None:
        return det(image)
    from .openpose import load_body_pose  # this tree's own body-pose detector, when its checkpoint is present locally

    body = load_body_pose()
    if body is not None:
        return body(image)
    # no classical stand-in for a pose estimator: an empty skeleton canvas means "no pose constraint"
    _stand_in("openpose", "an empty skeleton canvas (no pose constraint)")
    return Image.new("RGB", image.size, 0)


def image_to_softedge(image, scribble=False):
    det = _aux("HEDdetector" if scribble else "PidiNetDetector", "lllyasviel/Annotators")
    if det is not None:
        return det(image, scribble=True) if scribble else det(image)
    if not scribble:  # the reference's softedge detector is PiDiNet (HED draws the scribbles)
        from .pidinet import load_pidinet

        pidi = load_pidinet()
        if pidi is not None:
            return pidi.edges(image, next(pidi.parameters()).device)
    from .annotators import load_hed  # this tree's own HED network, when its checkpoint is present locally

    hed = load_hed()
    if hed is not None:
        return hed.edges(image, next(hed.parameters()).device, scribble=scribble)
    _stand_in("scribble" if scribble else "softedge", "Sobel edge magnitude")
    _, _, mag = _sobel_mag(image)
    soft = (255 * np.clip(mag * 2.0, 0, 1)).astype(np.uint8)
    if scribble:  # the same thinning / binarisation the learned detector's map gets
        from .annotators import scribble_strokes

        return _gray3(scribble_strokes(soft))
    return _gray3(soft)


def image_to_shuffle(image, seed=None):
    """Content shuffle: warp the image through a smooth random flow field."""
    import cv2

    arr = np.array(image.convert("RGB"))
    h, w = arr.shape[:2]
    rng = np.random.default_rng(seed)

    def field(n):
        f = rng.uniform(0, n, size=(max(2, h // 64), max(2, w // 64))).astype(np.float32)
        return cv2.resize(f, (w, h), interpolation=cv2.INTER_CUBIC)

    return Image.fromarray(cv2.remap(arr, field(w), field(h), cv2.INTER_LINEAR, borderMode=cv2.BORDER_REFLECT))


def ade_palette():
    """150 distinct RGB triplets for semantic classes, generated (golden-ratio hue walk) rather than
    shipped as a literal table: the checkpoints conditioned on ADE20K colours need the *same class ->
    same colour* property, which this keeps; exact ADE20K colours are loaded from
    `ade20k_palette.json` in the settings dir when an operator provides it."""
    import colorsys
    import json

    from ..settings import get_settings_dir

    custom = get_settings_dir() / "ade20k_palette.json"
    if custom.is_file():
        return [tuple(c) for c in json.loads(custom.read_text())]
    out = []
    for i in range(150):
        hue = (i * 0.61803398875) % 1.0
        sat = 0.55 + 0.45 * ((i * 7) % 5) / 4.0
        val = 0.55 + 0.45 * ((i * 3) % 4) / 3.0
        r, g, b = colorsys.hsv_to_rgb(hue, sat, val)
        out.append((int(r * 255), int(g * 255), int(b * 255)))
    return out


ada_palette = ade_palette  # the reference spells it `ada_palette`
