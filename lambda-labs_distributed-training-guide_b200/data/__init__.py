from .datasets import SyntheticTokens, TokenBlocks, group_texts, load_and_preprocess_data, resolve_seq_length
from .loader import default_data_collator, make_dataloader, set_epoch
from .packed import (EOS_ID_DEFAULT, block_diagonal_causal_mask, cu_seqlens_from_positions, packed_positions,
                     pad_and_collate)

__all__ = ["SyntheticTokens", "TokenBlocks", "group_texts", "load_and_preprocess_data", "resolve_seq_length",
           "default_data_collator", "make_dataloader", "set_epoch", "EOS_ID_DEFAULT",
           "block_diagonal_causal_mask", "cu_seqlens_from_positions", "packed_positions", "pad_and_collate"]
