"""img2txt through the job API with a LOCAL tiny BLIP checkpoint (reference: swarm/captioning/caption_image.py): the
processor / model classes named by the job's parameters are loaded once, stay resident, and the caption comes back as a
text artifact hashed over the raw string."""
import asyncio
import base64
import hashlib
import json

import pytest
import torch

from tests.helpers import png_data_uri


@pytest.fixture(autouse=True)
def _env(tmp_path, monkeypatch):
    monkeypatch.setenv("SDAAS_ROOT", str(tmp_path))
    monkeypatch.setenv("SDAAS_MODEL_ROOT", str(tmp_path / "models"))
    monkeypatch.setenv("SDAAS_LOCAL_INPUTS", "1")
    from chiaswarm_b200.utils.weight_cache import pipeline_cache

    pipeline_cache().clear()


def _tiny_blip(root):
    from transformers import (BertTokenizerFast, BlipConfig, BlipForConditionalGeneration, BlipImageProcessor, BlipProcessor)

    root.mkdir(parents=True)
    words = ["[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]", "[DEC]", "a", "picture", "of", "cat", "dog", "the", "on", "grass"]
    (root / "vocab.txt").write_text("\n".join(words) + "\n")
    tok = BertTokenizerFast(vocab_file=str(root / "vocab.txt"))
    tok.add_special_tokens({"bos_token": "[DEC]"})
    cfg = BlipConfig(
        text_config=dict(vocab_size=len(words), hidden_size=32, num_hidden_layers=2, num_attention_heads=2, intermediate_size=64,
                         encoder_hidden_size=32, max_position_embeddings=32, bos_token_id=5, pad_token_id=0, sep_token_id=3,
                         eos_token_id=3),
        vision_config=dict(hidden_size=32, intermediate_size=64, num_hidden_layers=2, num_attention_heads=2, image_size=32,
                           patch_size=8),
        projection_dim=16, image_text_hidden_size=16)
    torch.manual_seed(0)
    BlipForConditionalGeneration(cfg).save_pretrained(str(root))
    BlipProcessor(BlipImageProcessor(size={"height": 32, "width": 32}), tok).save_pretrained(str(root))


def _run(job):
    from chiaswarm_b200.generator import do_work
    from chiaswarm_b200.gpu.device import Device

    return asyncio.run(do_work(job, Device("cpu")))


def test_caption_job_with_a_local_checkpoint(tmp_path):
    from chiaswarm_b200.utils.weight_cache import pipeline_cache

    _tiny_blip(tmp_path / "models" / "Salesforce" / "blip-tiny")
    job = {"id": "c1", "workflow": "img2txt", "model_name": "Salesforce/blip-tiny", "start_image_uri": png_data_uri((40, 30)),
           "parameters": {"processor_type": "BlipProcessor", "model_type": "BlipForConditionalGeneration"}}
    hits0, misses0 = pipeline_cache().hits, pipeline_cache().misses          # (process-wide counters: other tests count too)
    r = _run(dict(job, parameters=dict(job["parameters"])))
    assert "error" not in r["pipeline_config"], r["pipeline_config"]
    caption = r["pipeline_config"]["caption"]
    art = r["artifacts"]["primary"]
    assert art["content_type"] == "application/json"
    assert json.loads(base64.b64decode(art["blob"])) == {"caption": caption}
    assert art["sha256_hash"] == hashlib.sha256(caption.encode()).hexdigest()     # the raw string is hashed (reference quirk)
    cache = pipeline_cache()
    assert len(cache) == 1 and cache.bytes_resident() > 10_000 and cache.misses == misses0 + 1
    # the second job (prompted captioning this time) finds processor and model resident and is deterministic
    r2 = _run(dict(job, id="c2", prompt="a picture of", parameters=dict(job["parameters"])))
    r3 = _run(dict(job, id="c3", prompt="a picture of", parameters=dict(job["parameters"])))
    assert "error" not in r2["pipeline_config"] and r2["pipeline_config"]["caption"] == r3["pipeline_config"]["caption"]
    assert cache.misses == misses0 + 1 and cache.hits == hits0 + 2
    # a class name transformers does not have is a caption-shaped error, not an exception
    r4 = _run(dict(job, id="c4", parameters={"processor_type": "NoSuchProcessor", "model_type": "BlipForConditionalGeneration"}))
    assert "NoSuchProcessor" in r4["pipeline_config"]["error"]


def test_cache_counts_whatever_modules_an_entry_holds():
    from chiaswarm_b200.utils.weight_cache import _nbytes

    lin = torch.nn.Linear(10, 10)                       # 110 floats
    bn = torch.nn.BatchNorm1d(4)                        # 8 parameters + 8 floats and one int64 of buffers

    class Holder:
        def __init__(self):
            self.unet, self.vocoder, self.note = lin, bn, "x"

    assert _nbytes(lin) == 110 * 4
    assert _nbytes((object(), lin)) == 110 * 4 and _nbytes({"a": lin, "b": lin}) == 110 * 4   # shared tensors count once
    assert _nbytes(Holder()) == 110 * 4 + 16 * 4 + 8
    assert _nbytes("nothing") == 0
