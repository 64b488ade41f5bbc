class DictConfig(dict):
    pass
class OmegaConf:
    @staticmethod
    def register_new_resolver(*a, **k):
        pass
