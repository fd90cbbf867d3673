from .state import (has_checkpoint, load_full, load_rng, load_state, new_state, save_full, save_rng, save_state,
                    set_determinism)

__all__ = ["has_checkpoint", "load_full", "load_rng", "load_state", "new_state", "save_full", "save_rng",
           "save_state", "set_determinism"]
