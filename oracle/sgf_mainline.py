"""A main-line SGF reader for the harness -- TEST INFRASTRUCTURE ONLY.

PyPI `sgf` (which AlphaGo/util.py:100-128 uses) is not installed here.  This
returns what util.sgf_iter_states needs from a file: board size, set-up stones
(AB/AW), the player to start (PL) and the main-line moves.  Coordinates follow
util.py:21-30: first letter = x, second = y; '' or 'tt' = pass.
"""
import re

_PROP = re.compile(r"([A-Z]+)((?:\s*\[(?:\\.|[^\]\\])*\])+)", re.S)
_VAL = re.compile(r"\[((?:\\.|[^\]\\])*)\]", re.S)
_LETTERS = "abcdefghijklmnopqrstuvwxyz"


def _xy(v):
    if v == "" or v == "tt":
        return None
    return (_LETTERS.index(v[0].lower()), _LETTERS.index(v[1].lower()))


def _nodes_mainline(text):
    """Splits the first game tree into nodes, following only the first variation."""
    depth, i, n, out, cur, skip_depth = 0, 0, len(text), [], None, None
    in_val = False
    while i < n:
        c = text[i]
        if in_val:
            if c == "\\":
                cur.append(text[i:i + 2]); i += 2; continue
            if c == "]":
                in_val = False
            cur.append(c)
        elif c == "[":
            in_val = True
            if cur is not None:
                cur.append(c)
        elif c == "(":
            depth += 1
        elif c == ")":
            depth -= 1
            if skip_depth is None:
                skip_depth = depth      # main line ended at this level; ignore later siblings
            if depth == 0:
                break
        elif c == ";":
            if skip_depth is not None and depth > skip_depth:
                # a sibling variation of a finished main line: skip it entirely
                cur = None
            else:
                cur = []
                out.append(cur)
        elif cur is not None:
            cur.append(c)
        i += 1
    return ["".join(x) for x in out]


def read(text):
    """-> dict(size, black=[(x,y)], white=[(x,y)], to_play=+1/-1, moves=[(color, (x,y)|None)])"""
    nodes = _nodes_mainline(text)
    game = {"size": 19, "black": [], "white": [], "to_play": +1, "moves": []}
    for k, node in enumerate(nodes):
        props = {m.group(1): _VAL.findall(m.group(2)) for m in _PROP.finditer(node)}
        if k == 0:
            if "SZ" in props:
                game["size"] = int(props["SZ"][0])
            game["black"] = [_xy(v) for v in props.get("AB", [])]
            game["white"] = [_xy(v) for v in props.get("AW", [])]
            if "PL" in props:
                game["to_play"] = +1 if props["PL"][0] == "B" else -1
        if "W" in props:
            game["moves"].append((-1, _xy(props["W"][0])))
        elif "B" in props:
            game["moves"].append((+1, _xy(props["B"][0])))
    return game
