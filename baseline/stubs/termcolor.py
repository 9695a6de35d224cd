"""Stand-in for termcolor (absent from this image): no colours."""


def colored(text, *a, **k):
    return text


def cprint(text, *a, **k):
    print(text)
