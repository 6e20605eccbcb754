def reset(value):
    """torch_geometric.nn.inits.reset: recurse into children, call reset_parameters."""
    if hasattr(value, "reset_parameters"):
        value.reset_parameters()
    else:
        for child in value.children() if hasattr(value, "children") else []:
            reset(child)
