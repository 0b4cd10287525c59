"""The encoder of the reference's one other model-bearing op, `map_summarize` (ops/map_summarize.py:17-59:
BART `generate`, beams 4), expressed as a .b2m model. TEST INFRASTRUCTURE.

A BART encoder layer is the layer this library runs — post-LayerNorm self-attention and FFN, erf-GELU,
heads of 64 — and only the embeddings differ: no token types, learned positions stored with an offset
of 2, LayerNorm after the sum (transformers' BartEncoder / BartLearnedPositionalEmbedding). So a
BartEncoder's weights map onto the model file one to one: q/k/v projections stacked into Wqkv,
`embed_positions.weight[2:]` as the position table, a zero token-type table. The pooler and classifier
tensors the file format also holds are filled with zeros; nothing of the encoder reads them.

What this does NOT give is `map_summarize`: that op is the decoder with beam search over downloaded
weights, none of which can be run or checked offline."""
import numpy as np
import torch


def random_bart(vocab_size=2048, layers=6, seed=7):
    from transformers import BartConfig, BartModel
    torch.manual_seed(seed)
    cfg = BartConfig(vocab_size=vocab_size, d_model=768, encoder_layers=layers, decoder_layers=1,
                     encoder_attention_heads=12, decoder_attention_heads=12, encoder_ffn_dim=3072,
                     decoder_ffn_dim=3072, max_position_embeddings=512, activation_function="gelu", dropout=0.0,
                     attention_dropout=0.0, activation_dropout=0.0, scale_embedding=False, pad_token_id=1)
    model = BartModel(cfg).eval()
    # random-init LayerNorms are the identity and biases zero: perturb them so every affine path counts
    g = torch.Generator().manual_seed(seed + 1)
    with torch.no_grad():
        for name, p in model.encoder.named_parameters():
            if "layer_norm" in name or "layernorm" in name:
                p.add_(0.1 * torch.randn(p.shape, generator=g))
            elif name.endswith(".bias"):
                p.add_(0.05 * torch.randn(p.shape, generator=g))
    return model


def encoder_as_model_file(model, path):
    """Write the BartEncoder of `model` as a kind-0 .b2m file; returns (path, ModelConfig)."""
    import model_file as mf
    enc = model.encoder
    sd = {k: v.detach().float().numpy() for k, v in enc.state_dict().items()}
    H = sd["embed_tokens.weight"].shape[1]
    L = len(enc.layers)
    cfg = mf.ModelConfig(hidden=H, layers=L, heads=H // 64, intermediate=sd["layers.0.fc1.weight"].shape[0],
                         vocab_size=sd["embed_tokens.weight"].shape[0],
                         max_positions=sd["embed_positions.weight"].shape[0] - 2, type_vocab=1, num_labels=2,
                         ln_eps=1e-5)
    T = {"word_emb": sd["embed_tokens.weight"],
         "pos_emb": sd["embed_positions.weight"][2:],          # BartLearnedPositionalEmbedding: offset 2
         "type_emb": np.zeros((1, H), np.float32),
         "emb_ln_g": sd["layernorm_embedding.weight"], "emb_ln_b": sd["layernorm_embedding.bias"],
         "Wp": np.zeros((H, H), np.float32), "bp": np.zeros(H, np.float32),
         "Wc": np.zeros((2, H), np.float32), "bc": np.zeros(2, np.float32)}
    for l in range(L):
        p, q = f"l{l}.", f"layers.{l}."
        T[p + "Wqkv"] = np.concatenate([sd[q + f"self_attn.{x}_proj.weight"] for x in "qkv"], 0)
        T[p + "bqkv"] = np.concatenate([sd[q + f"self_attn.{x}_proj.bias"] for x in "qkv"], 0)
        T[p + "Wo"], T[p + "bo"] = sd[q + "self_attn.out_proj.weight"], sd[q + "self_attn.out_proj.bias"]
        T[p + "ln1_g"], T[p + "ln1_b"] = sd[q + "self_attn_layer_norm.weight"], sd[q + "self_attn_layer_norm.bias"]
        T[p + "W1"], T[p + "b1"] = sd[q + "fc1.weight"], sd[q + "fc1.bias"]
        T[p + "W2"], T[p + "b2"] = sd[q + "fc2.weight"], sd[q + "fc2.bias"]
        T[p + "ln2_g"], T[p + "ln2_b"] = sd[q + "final_layer_norm.weight"], sd[q + "final_layer_norm.bias"]
    return mf.write_model_file(path, cfg, tensors=T), cfg


@torch.no_grad()
def bart_hidden_states(model, ids, mask):
    """[layers + 1, n, S, H] float32 from transformers' BartEncoder (hidden_states), pad rows zeroed."""
    out = model.encoder(input_ids=torch.from_numpy(ids).long(), attention_mask=torch.from_numpy(mask).long(),
                        output_hidden_states=True)
    hs = torch.stack(out.hidden_states).numpy().astype(np.float32)
    return hs * mask[None, :, :, None].astype(np.float32)
