class MultiAgentEnv:
    """The base class RLlib's multi-agent environments derive from: a constructor without arguments is all the reference's
    wrapper uses of it (rllib_envs/marl_wrapper.py:8-11)."""

    def __init__(self):
        pass
