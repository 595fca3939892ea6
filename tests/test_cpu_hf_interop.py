"""Hugging Face checkpoint directories load into the Llama / GPT-2 models here and give the loss `transformers` gives on the same weights
(the reference fine-tunes *pretrained* Llama-2: 02_development/distributed_utils.py:462-467); exports load back into `transformers`.
Oracle: the stock `transformers` implementations, fp32, CPU."""
import json

import pytest
import torch


def _hf_llama(tmp_path, kv_heads=2):
    from transformers import LlamaConfig as HC, LlamaForCausalLM as HM

    torch.manual_seed(0)
    hf = HM(HC(vocab_size=320, hidden_size=64, intermediate_size=176, num_hidden_layers=2, num_attention_heads=4,
               num_key_value_heads=kv_heads, rms_norm_eps=1e-5, rope_theta=10000.0, tie_word_embeddings=False, pad_token_id=2,
               max_position_embeddings=64)).eval()
    hf.save_pretrained(tmp_path / "llama", safe_serialization=True)
    return hf, tmp_path / "llama"


def _batch(vocab):
    g = torch.Generator().manual_seed(3)
    ids = torch.randint(3, vocab, (2, 24), generator=g)
    mask = torch.ones_like(ids)
    mask[1, 15:] = 0
    labels = ids.clone()
    labels[mask == 0] = -100
    return ids, mask, labels


@pytest.mark.parametrize("kv_heads", [4, 2])          # multi-head and grouped-query checkpoints
def test_llama_checkpoint_dir_gives_the_transformers_loss(tmp_path, kv_heads):
    from hyperion_b200.models import hf as H

    ref, path = _hf_llama(tmp_path, kv_heads)
    assert H.is_checkpoint_dir(path) and not H.is_checkpoint_dir("NousResearch/Llama-2-7b-hf") and not H.is_checkpoint_dir(None)
    model = H.llama_from_pretrained(path).eval()
    assert model.config.num_key_value_heads == kv_heads and model.config.pad_token_id == 2
    ids, mask, labels = _batch(320)
    with torch.no_grad():
        want = ref(input_ids=ids, attention_mask=mask, labels=labels)
        got_loss = model(input_ids=ids, attention_mask=mask, labels=labels).loss
        got_logits = model(input_ids=ids, attention_mask=mask).logits
    assert float(got_loss) == pytest.approx(float(want.loss), rel=1e-5)
    torch.testing.assert_close(got_logits[mask.bool()], want.logits[mask.bool()], rtol=1e-4, atol=1e-4)
    # gradients too: one backward through both, same dW for a projection and the embedding
    ref.train(); model.train()
    ref(input_ids=ids, attention_mask=mask, labels=labels).loss.backward()
    model(input_ids=ids, attention_mask=mask, labels=labels).loss.backward()
    for name in ("model.layers.1.self_attn.k_proj.weight", "model.layers.0.mlp.down_proj.weight", "model.embed_tokens.weight", "lm_head.weight"):
        torch.testing.assert_close(dict(model.named_parameters())[name].grad, dict(ref.named_parameters())[name].grad, rtol=2e-3, atol=1e-6)


def test_llama_export_loads_back_into_transformers(tmp_path):
    from transformers import AutoModelForCausalLM

    from hyperion_b200.models import LlamaConfig, LlamaForCausalLM
    from hyperion_b200.models import hf as H

    torch.manual_seed(1)
    ours = LlamaForCausalLM(LlamaConfig.tiny()).eval()
    out = H.save_hf(ours, tmp_path / "export")
    assert json.loads((out / "config.json").read_text())["model_type"] == "llama"
    back = AutoModelForCausalLM.from_pretrained(out).eval()
    ids, mask, labels = _batch(512)
    with torch.no_grad():
        a = float(ours(input_ids=ids, attention_mask=mask, labels=labels).loss)
        b = float(back(input_ids=ids, attention_mask=mask, labels=labels).loss)
    assert a == pytest.approx(b, rel=1e-5)
    again = H.llama_from_pretrained(out)
    for (k, v), (k2, v2) in zip(ours.state_dict().items(), again.state_dict().items()):
        assert k == k2 and torch.equal(v, v2)


def test_gpt2_checkpoint_dir_round_trip(tmp_path):
    from transformers import AutoModelForCausalLM, GPT2Config as HC, GPT2LMHeadModel as HM

    from hyperion_b200.models import hf as H

    torch.manual_seed(0)
    ref = HM(HC(vocab_size=300, n_positions=32, n_embd=48, n_layer=2, n_head=3, resid_pdrop=0.0, embd_pdrop=0.0, attn_pdrop=0.0)).eval()
    ref.save_pretrained(tmp_path / "gpt2", safe_serialization=True)
    model = H.gpt2_from_pretrained(tmp_path / "gpt2").eval()
    assert model.config.n_head == 3 and model.h[0].attn.c_attn.weight.shape == (144, 48)       # [out, in], HF's Conv1D is [in, out]
    ids = torch.randint(0, 300, (2, 20), generator=torch.Generator().manual_seed(5))
    with torch.no_grad():
        want = ref(input_ids=ids, labels=ids)
        got = model.loss(ids)
        logits = model(ids)
    assert float(got) == pytest.approx(float(want.loss), rel=1e-5)
    torch.testing.assert_close(logits, want.logits, rtol=1e-4, atol=1e-4)
    out = H.save_hf(model, tmp_path / "export")
    back = AutoModelForCausalLM.from_pretrained(out).eval()
    with torch.no_grad():
        assert float(back(input_ids=ids, labels=ids).loss) == pytest.approx(float(want.loss), rel=1e-5)


def test_sharded_safetensors_and_bin_checkpoints(tmp_path):
    """The 7B checkpoint comes as several shards; older ones as pytorch_model*.bin."""
    from transformers import LlamaConfig as HC, LlamaForCausalLM as HM

    from hyperion_b200.models import hf as H

    torch.manual_seed(2)
    ref = HM(HC(vocab_size=320, hidden_size=64, intermediate_size=176, num_hidden_layers=3, num_attention_heads=4, num_key_value_heads=4,
                tie_word_embeddings=False, pad_token_id=2)).eval()
    ref.save_pretrained(tmp_path / "sharded", safe_serialization=True, max_shard_size="100KB")
    assert len(list((tmp_path / "sharded").glob("*.safetensors"))) > 2
    (tmp_path / "bin").mkdir()
    (tmp_path / "bin/config.json").write_text((tmp_path / "sharded/config.json").read_text())
    torch.save(ref.state_dict(), tmp_path / "bin/pytorch_model.bin")
    want = {k: v for k, v in ref.state_dict().items()}
    for d in ("sharded", "bin"):
        got = H.llama_from_pretrained(tmp_path / d, dtype=torch.bfloat16).state_dict()       # converted while copying
        assert set(got) == set(want)
        for k in want:
            assert got[k].dtype == torch.bfloat16 and torch.equal(got[k], want[k].to(torch.bfloat16)), k


def test_unsupported_checkpoints_are_refused(tmp_path):
    from hyperion_b200.models import hf as H

    base = dict(vocab_size=32, hidden_size=16, intermediate_size=32, num_hidden_layers=1, num_attention_heads=2)
    with pytest.raises(NotImplementedError, match="rope_scaling"):
        H.llama_config_from_hf({**base, "rope_scaling": {"rope_type": "llama3", "factor": 8.0}})
    with pytest.raises(NotImplementedError, match="bias"):
        H.llama_config_from_hf({**base, "attention_bias": True})
    with pytest.raises(NotImplementedError, match="tie_word_embeddings"):
        H.llama_config_from_hf({**base, "tie_word_embeddings": True})
    with pytest.raises(NotImplementedError, match="gelu_new"):
        H.gpt2_config_from_hf({"vocab_size": 8, "n_embd": 8, "n_layer": 1, "n_head": 1, "activation_function": "relu"})
    assert H.llama_config_from_hf({**base, "eos_token_id": 7}).pad_token_id == 7         # pad = eos, like the reference
    (tmp_path / "empty").mkdir()
    (tmp_path / "empty/config.json").write_text(json.dumps({**base, "model_type": "llama"}))
    with pytest.raises(FileNotFoundError):
        H.llama_from_pretrained(tmp_path / "empty")


def test_llama_trainer_starts_from_a_checkpoint_dir(tmp_path, capsys):
    import os

    from hyperion_b200.parallel.runtime import cleanup
    from hyperion_b200.train import train_llama_fsdp

    ref, path = _hf_llama(tmp_path)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT="29533", RANK="0", WORLD_SIZE="1")
    try:
        hist = train_llama_fsdp(0, 1, epochs=1, base_dir=str(tmp_path), model_id=str(path), lora=True, batch_size=2, backend="gloo",
                                seq_len=16, steps_per_epoch=2, save=False, graph=False)
    finally:
        cleanup()
    out = capsys.readouterr().out
    assert "weights read from" in out and "LoRA adapters" in out and hist
