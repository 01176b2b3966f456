import re
import textwrap


def a10n(string):
    return string if len(string) < 3 else "%s%s%s" % (string[0], len(string) - 2, string[-1])


def asciify(text, ignore=False):
    return text.encode("ascii", "ignore" if ignore else "replace")


def format_int_list(int_list, delim=",", range_delim="-", delim_space=False):
    return (delim + (" " if delim_space else "")).join(str(i) for i in int_list)


def parse_int_list(range_string, delim=",", range_delim="-"):
    out = []
    for part in range_string.split(delim):
        part = part.strip()
        if not part:
            continue
        if range_delim in part[1:]:
            a, b = part.split(range_delim, 1)
            out.extend(range(int(a), int(b) + 1))
        else:
            out.append(int(part))
    return sorted(out)


def iter_splitlines(text):
    yield from text.splitlines()


def slugify(text, delim="_", lower=True, ascii=False):
    s = re.sub(r"[^\w]+", delim, text).strip(delim)
    return s.lower() if lower else s


def unwrap_text(text, ending="\n\n"):
    return ending.join(" ".join(p.split()) for p in textwrap.dedent(text).split("\n\n"))
