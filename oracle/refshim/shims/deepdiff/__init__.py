class DeepDiff(dict):
    def __init__(self, a, b, *args, **kwargs):
        super().__init__()
        if a != b:
            self["values_changed"] = {"root": {"old_value": a, "new_value": b}}
