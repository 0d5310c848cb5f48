"""`add_enhance_arguments` with the reference's contract
(open_universe/inference_utils/signature_to_parser.py:26-66): one `--<kwarg>` option per
type-hinted keyword of `model.enhance`, typed by the first type argument of the hint,
defaulting to `model.diff_kwargs[<kwarg>]` when present."""
import typing


def add_enhance_arguments(model, parser):
    enhance = getattr(model, "enhance", None)
    if not callable(enhance):
        raise ValueError("Model does not have an `enhance` method.")
    hints = typing.get_type_hints(enhance)
    hints.pop("return", None)
    defaults = getattr(model, "diff_kwargs", {})
    group = parser.add_argument_group("enhance", "Arguments of enhance function")
    for name, hint in hints.items():
        inner = typing.get_args(hint)
        caster = inner[0] if inner else hint
        group.add_argument(f"--{name}", default=defaults.get(name, None), type=caster)
    return parser
