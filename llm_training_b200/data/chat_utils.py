"""Chat-template tokenisation shared by instruction / preference tuning: token ids + assistant-only labels."""
from typing import Any


def encode_conversations(tokenizer: Any, conversations: list[list[dict[str, str]]], chat_template: str | None
                         ) -> tuple[list[list[int]], list[list[int]]]:
    """-> (input_ids, labels) where labels are -100 outside ``{% generation %}`` spans."""
    enc = tokenizer.apply_chat_template(
        conversations, chat_template=chat_template, tokenize=True, return_dict=True,
        return_assistant_tokens_mask=True, tokenizer_kwargs=dict(return_attention_mask=False, verbose=False))
    ids_rows, label_rows = [], []
    for ids, mask in zip(enc['input_ids'], enc['assistant_masks']):
        ids = list(ids)
        ids_rows.append(ids)
        label_rows.append([t if m == 1 else -100 for t, m in zip(ids, mask)])
    return ids_rows, label_rows


def pad_rows(rows: list[list[int]], value: int, left: bool, multiple_of: int | None):
    import torch
    n = max(len(r) for r in rows)
    if multiple_of is not None:
        n = ((n // multiple_of) + 1) * multiple_of      # always at least one pad column, like the reference
    out = torch.full((len(rows), n), value, dtype=torch.long)
    for i, r in enumerate(rows):
        if r:
            t = torch.as_tensor(r, dtype=torch.long)
            if left:
                out[i, n - len(r):] = t
            else:
                out[i, :len(r)] = t
    return out
