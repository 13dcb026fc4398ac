"""Drop-in entry point: `python prepare_weights.py <bin_dir> <new_file_dir>` (same CLI and output layout
as the reference's prepare_weights.py).  Implementation: flexible-llm-sharding_b200/prepare_weights.py."""
from flexible_llm_sharding_b200.prepare_weights import main, split_into_layers  # noqa: F401

if __name__ == "__main__":
    main()
