"""Mirror of the reference's ``model`` package: ``model.net`` and ``model.graph_met_network``."""
