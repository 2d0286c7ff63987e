class DegreeScalerAggregation:      # named by GN0/torch_script_models.py's imports only
    def __init__(self, *a, **k):
        raise NotImplementedError("pyg_stub: DegreeScalerAggregation is not on the pinned path")
