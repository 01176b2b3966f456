def get_include_dir():
    return "/nonexistent"
