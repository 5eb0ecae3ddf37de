"""Small helpers used by the side tools.  Ref: src/aux_modules/utils/misc.py:1-288 (coloured prints,
keyboard prompts, Mallet corpus <-> DataFrame conversion, second config reader)."""
from __future__ import annotations

from typing import List, Optional

import pandas as pd

from ..utils.auxiliary_functions import read_config_experiments  # noqa: F401  (same reader, re-exported)

_COLORS = {"red": "\033[91m", "green": "\033[92m", "yellow": "\033[93m", "blue": "\033[94m", "end": "\033[0m"}


def _cprint(color: str, *args) -> None:
    print(_COLORS[color] + " ".join(str(a) for a in args) + _COLORS["end"])


def printgr(*args) -> None:
    _cprint("green", *args)


def printred(*args) -> None:
    _cprint("red", *args)


def printmag(*args) -> None:
    _cprint("blue", *args)


def is_integer(user_str: str) -> bool:
    try:
        int(user_str)
        return True
    except (TypeError, ValueError):
        return False


def is_float(user_str: str) -> bool:
    try:
        float(user_str)
        return True
    except (TypeError, ValueError):
        return False


def var_num_keyboard(vartype: str, default, question: str, reader=input):
    """Ask for an int/float on the keyboard; empty input returns ``default``."""
    ans = reader(f"{question} [default: {default}]: ").strip()
    if not ans:
        return default
    if vartype == "int" and is_integer(ans):
        return int(ans)
    if vartype == "float" and is_float(ans):
        return float(ans)
    return default


def request_confirmation(msg: str = "     Are you sure?", reader=input) -> bool:
    return reader(msg + " (y/n): ").strip().lower() in ("y", "yes")


def query_options(options: List[str], msg: Optional[str] = None, reader=input) -> int:
    """Prints numbered options and returns the chosen index."""
    if msg:
        print(msg)
    for i, o in enumerate(options):
        print(f" {i}. {o}")
    while True:
        ans = reader("Select option: ").strip()
        if is_integer(ans) and 0 <= int(ans) < len(options):
            return int(ans)


def mallet_corpus_to_df(corpus_file: str) -> pd.DataFrame:
    """Mallet text format (``<id> <label> tok tok ...`` per line) -> DataFrame[id, text]."""
    ids, texts = [], []
    with open(corpus_file, encoding="utf-8") as f:
        for line in f:
            parts = line.rstrip("\n").split(" ", 2)
            if len(parts) == 3:
                ids.append(parts[0])
                texts.append(parts[2])
    return pd.DataFrame({"id": ids, "text": texts})


def corpus_df_to_mallet(corpus_df: pd.DataFrame, out_file: str) -> None:
    with open(out_file, "w", encoding="utf-8") as f:
        for _id, text in zip(corpus_df["id"], corpus_df["text"]):
            f.write(f"{_id} 0 {text}\n")


# ---------------------------------------------------------------------------------------------
# Remaining helpers of src/aux_modules/utils/misc.py (keyboard string reader, title clean-up, pickles)
# ---------------------------------------------------------------------------------------------
def var_string_keyboard(option: str, default, question: str, reader=input):
    """Ask for a structured value on the keyboard (ref: misc.py:86-136).  ``option`` selects the expected
    format: ``comma_separated`` (a parenthesised tuple of ints such as ``(50, 50)``), ``bool``
    (``True`` / ``False``), ``dict`` (a JSON object) or ``str``.  Empty or malformed input returns
    ``default`` (with a notice when something was typed)."""
    import json

    raw = reader(f"{question} [{default}]: ")
    invalid = "The value you provided is not valid. Using default value."
    value = None
    if option == "comma_separated":
        parts = [p.strip() for p in raw.split(",") if p]
        if len(parts) <= 1:
            print(invalid)
            return default
        try:
            value = tuple(int(p) for p in raw.strip()[1:-1].split(","))
        except ValueError:
            print(invalid)
            return default
    elif option == "bool":
        if raw not in ("True", "False"):
            print(invalid)
            return default
        value = raw == "True"
    elif option == "dict":
        if raw in ("", "None"):
            return default
        try:
            value = json.loads(raw)
        except ValueError:
            print(invalid)
            return default
    elif option == "str":
        value = raw if raw != "" else None
    if value is None:
        if raw != "":
            print(invalid)
        return default
    return value


def format_title(tgt_str: str) -> str:
    """Title as stored in the corpus tables: double quotes and line breaks removed (misc.py:185-190)."""
    return tgt_str.replace('"', "").replace("\n", "")


def unpickler(file: str):
    import pickle

    with open(file, "rb") as f:
        return pickle.load(f)


def pickler(file: str, ob) -> None:
    import pickle

    with open(file, "wb") as f:
        pickle.dump(ob, f)
