"""Drop-in for the Bayesian-optimisation part of ``polyaxon/hpsearch`` (reference layout kept:
``hpsearch.search_managers.bayesian_optimization.{manager,optimizer,acquisition_function,space}``,
``hpsearch.search_managers.{base,utils}``, ``hpsearch.schemas.bayesian_optimization``)."""
