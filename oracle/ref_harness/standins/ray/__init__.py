"""Stand-in for `ray` (not installed here).  TEST INFRASTRUCTURE ONLY: just enough for the reference's
`rllib_envs/marl_wrapper.py` to import -- the wrapper only subclasses `ray.rllib.env.MultiAgentEnv`."""
__version__ = "stand-in"
