"""Minimal stand-in for the `configargparse` package (not installed in this image), covering what
the reference's option surface uses (modules/model/utils/parser.py:9-57): `is_config_file`
arguments (`-c file`), `key = value` config lines feeding the same options as the command line
(command line wins), `store_true` flags written as `key=True`, unknown config keys passed
through to `parse_known_args`, string argument lists, and `_config_file_parser.serialize`.
"""
import argparse
import shlex

ArgumentDefaultsHelpFormatter = argparse.ArgumentDefaultsHelpFormatter
Namespace = argparse.Namespace

_TRUE = {"true", "yes", "1", "on"}
_FALSE = {"false", "no", "0", "off", "none", ""}


class _ConfigFileParser:
    @staticmethod
    def parse(stream):
        items = []
        for raw in stream:
            line = raw.strip()
            if not line or line[0] in "#;[":
                continue
            if "=" in line:
                key, value = line.split("=", 1)
            elif ":" in line:
                key, value = line.split(":", 1)
            else:
                key, value = line, "true"
            items.append((key.strip().lstrip("-"), value.strip()))
        return items

    @staticmethod
    def serialize(items):
        return "".join(f"{k} = {v}\n" for k, v in items.items())


class ArgumentParser(argparse.ArgumentParser):
    def __init__(self, *args, **kwargs):
        kwargs.pop("default_config_files", None)
        kwargs.pop("config_file_parser_class", None)
        super().__init__(*args, **kwargs)
        self._config_file_parser = _ConfigFileParser()
        self._config_dests = []

    def add_argument(self, *args, **kwargs):
        is_cfg = kwargs.pop("is_config_file", False) or kwargs.pop("is_config_file_arg", False)
        kwargs.pop("env_var", None)
        action = super().add_argument(*args, **kwargs)
        if is_cfg:
            self._config_dests.append(action)
        return action

    add = add_argument

    # ------------------------------------------------------------------------------------
    def _option_map(self):
        m = {}
        for a in self._actions:
            for s in a.option_strings:
                m[s.lstrip("-")] = a
        return m

    def _config_paths(self, argv):
        """Values given on the command line for the is_config_file options."""
        names = {s for a in self._config_dests for s in a.option_strings}
        paths, i = [], 0
        while i < len(argv):
            tok = argv[i]
            if tok in names and i + 1 < len(argv):
                paths.append(argv[i + 1])
                i += 2
                continue
            for n in names:
                if tok.startswith(n + "="):
                    paths.append(tok[len(n) + 1:])
            i += 1
        return paths

    def _config_args(self, argv):
        opts = self._option_map()
        extra = []
        for path in self._config_paths(argv):
            with open(path) as f:
                for key, value in self._config_file_parser.parse(f):
                    action = opts.get(key)
                    if action is None:
                        extra.append(f"--{key}={value}")  # unknown here: surfaces in parse_known_args
                    elif isinstance(action, (argparse._StoreTrueAction, argparse._StoreFalseAction)):
                        low = value.lower()
                        if low in _TRUE:
                            extra.append(action.option_strings[-1])
                        elif low not in _FALSE:
                            self.error(f"config {path}: {key} expects a boolean, got {value!r}")
                    elif action in self._config_dests:
                        continue
                    else:
                        extra += [action.option_strings[-1], value]
        return extra

    def parse_known_args(self, args=None, namespace=None):
        import sys
        if args is None:
            args = sys.argv[1:]
        elif isinstance(args, str):
            args = shlex.split(args)
        else:
            args = list(args)
        # config-file values first, so that explicit command-line values override them
        return super().parse_known_args(self._config_args(args) + args, namespace)

    def parse_args(self, args=None, namespace=None):
        ns, unknown = self.parse_known_args(args, namespace)
        if unknown:
            self.error("unrecognized arguments: %s" % " ".join(unknown))
        return ns


ArgParser = ArgumentParser
