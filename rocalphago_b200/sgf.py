"""Main line of the first game of an SGF collection -- what util.sgf_iter_states (util.py:100-128)
takes from the PyPI `sgf` parser, which this image does not have.

    game = read(text)
    game.size, game.add_black, game.add_white, game.first_player, game.moves  # [(color, (x, y) | None)]

Coordinates as util.py:21-30: first letter = x (column), second = y (row); '' and 'tt' are passes.
Only the root node's SZ / AB / AW / PL and each later node's first W (else B) value are used, as in
the reference; variations after the first are skipped.
"""
import string

BLACK, WHITE = 1, -1


class ParseException(ValueError):
    pass


class Game(object):
    def __init__(self, size, add_black, add_white, first_player, moves):
        self.size, self.add_black, self.add_white = size, add_black, add_white
        self.first_player, self.moves = first_player, moves


def _point(value):
    if value == "" or value == "tt":
        return None
    if len(value) < 2 or value[0].lower() not in string.ascii_lowercase or value[1].lower() not in string.ascii_lowercase:
        raise ParseException("bad point %r" % value)
    return (string.ascii_lowercase.index(value[0].lower()), string.ascii_lowercase.index(value[1].lower()))


def _main_line_nodes(text):
    """[{property: [values]}] for the nodes along the first variation of the first game tree."""
    i, n = 0, len(text)
    while i < n and text[i] != "(":
        i += 1
    if i == n:
        raise ParseException("no game tree")
    nodes = []
    props, ident, fresh = None, "", True
    while i < n:
        c = text[i]
        if c == ")":
            # Always descending into the first child tree, the first ')' outside a value closes the
            # deepest first variation: the main line ends here.
            return nodes
        elif c == ";":
            props, ident, fresh = {}, "", True
            nodes.append(props)
        elif c == "[":
            if props is None or not ident:
                raise ParseException("property value without an identifier at offset %d" % i)
            j, val = i + 1, []
            while j < n and text[j] != "]":
                if text[j] == "\\" and j + 1 < n:
                    j += 1
                val.append(text[j])
                j += 1
            if j >= n:
                raise ParseException("unterminated property value")
            props.setdefault(ident, []).append("".join(val))
            i, fresh = j, True
        elif c.isalpha():
            if fresh:                                    # first letter after a value or a ';'
                ident, fresh = "", False
            if c.isupper():                              # FF[3] allowed lower-case filler letters: ignored
                ident += c
        i += 1
    raise ParseException("unbalanced parentheses")


def read(text):
    nodes = _main_line_nodes(text)
    if not nodes:
        raise ParseException("empty game")
    root = nodes[0]
    size = int(root.get("SZ", ["19"])[0])
    add_black = [_point(v) for v in root.get("AB", [])]
    add_white = [_point(v) for v in root.get("AW", [])]
    first_player = BLACK if root.get("PL", ["B"])[0] == "B" else WHITE
    moves = []
    for node in nodes[1:]:
        # util.py:115-121: W wins over B when both are present; a node with neither repeats the
        # previous (move, player) in the reference -- such nodes do not occur in game records and
        # are skipped here.
        if "W" in node:
            moves.append((WHITE, _point(node["W"][0])))
        elif "B" in node:
            moves.append((BLACK, _point(node["B"][0])))
    return Game(size, add_black, add_white, first_player, moves)
