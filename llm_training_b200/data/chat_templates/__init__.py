"""Built-in chat templates whose assistant turns are wrapped in ``{% generation %}`` so that
``tokenizer.apply_chat_template(..., return_assistant_tokens_mask=True)`` yields the label mask
(reference: src/llm_training/data/chat_templates/__init__.py:6-37).  Lookup order of ``get_chat_template``:
an existing file path, then a built-in name, then the string itself as a literal template."""
import logging
from pathlib import Path

logger = logging.getLogger(__name__)

_DIR = Path(__file__).parent


class _BuiltinTemplates:
    def names(self) -> list[str]:
        return sorted(p.name[:-3] for p in _DIR.glob('*.j2'))

    def __contains__(self, name: str) -> bool:
        return isinstance(name, str) and '/' not in name and (_DIR / f'{name}.j2').is_file()

    def __getitem__(self, name: str) -> str:
        if name not in self:
            raise KeyError(f'Template `{name}` is not found; built-ins: {self.names()}')
        return (_DIR / f'{name}.j2').read_text()


CHAT_TEMPLATES = _BuiltinTemplates()


def get_chat_template(chat_template: str) -> str:
    try:
        is_file = len(chat_template) < 1024 and Path(chat_template).is_file()
    except OSError:
        is_file = False
    if is_file:
        logger.info(f'Found template file at `{chat_template}`.')
        return Path(chat_template).read_text()
    if chat_template in CHAT_TEMPLATES:
        logger.info(f'Using pre-defined chat template `{chat_template}`.')
        return CHAT_TEMPLATES[chat_template]
    logger.warning('`chat_template` is being used directly as a chat template. If this is not what you expected, '
                   f'use a file path or one of the built-in names: {CHAT_TEMPLATES.names()}')
    return chat_template
