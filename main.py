"""Drop-in entry point: `python main.py --model_path ... --prompt_pickle ... --output_file ...`
(same CLI as the reference's main.py).  The implementation lives in flexible-llm-sharding_b200/main.py."""
from flexible_llm_sharding_b200.main import main, run_sharded_llama  # noqa: F401

if __name__ == "__main__":
    main()
