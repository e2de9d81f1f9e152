"""Offline fixtures: a tiny word-level tokenizer (no network) and small model configs."""
import torch


def tiny_tokenizer(extra_tokens: tuple[str, ...] = ()):
    from tokenizers import Tokenizer, models, pre_tokenizers
    from transformers import PreTrainedTokenizerFast
    words = ['<pad>', '<s>', '</s>', '<unk>', '<|im_start|>', '<|im_end|>', 'system', 'user', 'assistant', '\n',
             *extra_tokens] + [f'w{i}' for i in range(200)] + list('abcdefghijklmnopqrstuvwxyz .,!?')
    vocab = {w: i for i, w in enumerate(dict.fromkeys(words))}
    tok = Tokenizer(models.WordLevel(vocab, unk_token='<unk>'))
    tok.pre_tokenizer = pre_tokenizers.Split(
        __import__('tokenizers').Regex(r'<\|im_start\|>|<\|im_end\|>|\n|[A-Za-z0-9]+|\S| '), behavior='isolated')
    fast = PreTrainedTokenizerFast(tokenizer_object=tok, bos_token='<s>', eos_token='</s>', pad_token='<pad>',
                                   unk_token='<unk>')
    fast.padding_side = 'right'
    return fast


TINY_LLAMA = dict(vocab_size=256, hidden_size=64, intermediate_size=128, num_hidden_layers=2, num_attention_heads=4,
                  num_key_value_heads=2, attn_implementation='sdpa')
