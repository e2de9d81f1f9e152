"""Machine-checked API parity with the reference source tree (read with ``ast``, never imported): every public module
path, class, constructor argument, public method and module-level function of ``/root/reference/src/llm_training`` must
exist under the same dotted path in this repository's ``llm_training`` alias package.  The allow-list below is the
complete set of deliberate omissions."""
import ast
import importlib
import inspect
import os

import pytest

REF = '/root/reference/src/llm_training'

ALLOWED_MISSING = {
    # private helpers of the reference
    'llm_training.lightning.callbacks.output_redirection._StreamRedirector',
    'llm_training.data.chat_templates._ChatTemplates',
    'llm_training.lms.orpo.orpo._ForwardOutput',
    'llm_training.lms.dpo.dpo._ForwardOutput',
    # vendored copy of transformers' AttentionMaskConverter; `prepare_4d_causal_attention_mask` is implemented directly
    'llm_training.ops.attention_op.AttentionMaskConverter',
    # a bool constructor argument of the same name exists here (the method is `_redirect_loggers`)
    'llm_training.lightning.callbacks.output_redirection.OutputRedirection.redirect_loggers',
}


def _is_validator(node: ast.FunctionDef) -> bool:
    """pydantic ``@field_validator`` / ``@model_validator`` methods are implementation details, not API."""
    for d in node.decorator_list:
        name = d.func if isinstance(d, ast.Call) else d
        text = ast.unparse(name)
        if 'validator' in text:
            return True
    return False


@pytest.mark.skipif(not os.path.isdir(REF), reason='reference source tree not present')
def test_public_api_of_the_reference_exists_here():
    import llm_training  # noqa: F401  (the alias package)
    problems: list[str] = []
    for d, _, files in os.walk(REF):
        for f in sorted(files):
            if not f.endswith('.py'):
                continue
            path = os.path.join(d, f)
            rel = os.path.relpath(path, REF)[:-3].replace(os.sep, '.')
            mod_name = 'llm_training' if rel == '__init__' else 'llm_training.' + rel.removesuffix('.__init__')
            tree = ast.parse(open(path).read())
            classes = [n for n in tree.body if isinstance(n, ast.ClassDef)]
            funcs = [n for n in tree.body if isinstance(n, ast.FunctionDef) and not n.name.startswith('_')]
            if not classes and not funcs:
                continue
            try:
                mod = importlib.import_module(mod_name)
            except Exception as e:  # noqa: BLE001
                problems.append(f'module {mod_name} does not import: {type(e).__name__}: {e}')
                continue
            for fn in funcs:
                if not hasattr(mod, fn.name) and f'{mod_name}.{fn.name}' not in ALLOWED_MISSING:
                    problems.append(f'function {mod_name}.{fn.name}')
            for c in classes:
                full = f'{mod_name}.{c.name}'
                ours = getattr(mod, c.name, None)
                if ours is None:
                    if full not in ALLOWED_MISSING:
                        problems.append(f'class {full}')
                    continue
                for n in c.body:
                    if not isinstance(n, ast.FunctionDef):
                        continue
                    if n.name == '__init__':
                        try:
                            sig = inspect.signature(ours.__init__)
                        except (TypeError, ValueError):
                            continue
                        if any(p.kind is p.VAR_KEYWORD for p in sig.parameters.values()):
                            continue
                        for a in n.args.args[1:] + n.args.kwonlyargs:
                            if a.arg not in sig.parameters:
                                problems.append(f'argument {full}.__init__({a.arg})')
                    elif not n.name.startswith('_') and not _is_validator(n) and not hasattr(ours, n.name) \
                            and f'{full}.{n.name}' not in ALLOWED_MISSING:
                        problems.append(f'method {full}.{n.name}')
    assert not problems, 'missing from this repository:\n' + '\n'.join(problems)
