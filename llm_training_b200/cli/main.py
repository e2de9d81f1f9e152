"""Console entry point ``llm-training`` (reference: src/llm_training/cli/main.py:4-5, pyproject.toml:31-32)."""


def main() -> None:
    from llm_training_b200.lightning.cli import LightningCLI
    LightningCLI()


if __name__ == '__main__':
    main()
