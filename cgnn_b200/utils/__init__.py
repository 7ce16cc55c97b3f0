from .Settings import SETTINGS
