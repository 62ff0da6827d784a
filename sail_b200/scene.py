"""Scene (mirror of /root/reference/src/scene/scene.js:11-119) and the reference scenes of
BASELINE.json / SURVEY.md 8(d) (``reference_scene(config)``)."""
from .camera import Camera
from .generator import (FILTER_PLUGINS, LIGHT_PLUGINS, MATERIAL_PLUGINS, SHAPE_PLUGINS, TEXTURE_PLUGINS, TRACE_PLUGINS,
                        PluginParams, query)
from .geometry import Object3D
from .light import GeometryLight, Light


class Scene:
    def __init__(self):
        self.camera = None
        self.objects = []
        self.lights = []
        self.sampleCount = 0
        self._trace = PluginParams("path")
        self._filter = PluginParams("color")
        self.select = None
        self.moving = False

    # scene.js:24-38: assigning a plugin *name* keeps the PluginParams object (and its params)
    @property
    def filter(self):
        return self._filter

    @filter.setter
    def filter(self, plugin):
        if query(FILTER_PLUGINS, plugin):
            self._filter.name = plugin

    @property
    def trace(self):
        return self._trace

    @trace.setter
    def trace(self, plugin):
        if query(TRACE_PLUGINS, plugin):
            self._trace.name = plugin

    @property
    def mat(self):
        """scene.js:40-42: projection x modelview (float64, row-major)."""
        return self.camera.projection @ self.camera.modelview

    @property
    def eye(self):
        return self.camera.eye

    def add(self, something):
        """scene.js:52-63."""
        if isinstance(something, Camera):
            self.camera = something
        elif isinstance(something, Object3D):
            self.objects.append(something)
        elif isinstance(something, Light):
            if isinstance(something, GeometryLight):
                self.objects.append(something.getGeometry(len(self.objects)))
            self.lights.append(something)

    def update(self):
        """scene.js:65-68 (the reference resets a *global* ``scene``; fixed to ``self``, Q13)."""
        self.camera.update()
        self.sampleCount = 0

    def tracerConfig(self):
        """scene.js:70-112: plugin discovery in first-seen order."""
        plugins = {"shape": [], "light": [], "material": [], "texture": [], "trace": self.trace}
        seen = {"shape": [], "light": [], "material": [], "texture": []}

        def note(kind, name):
            if name and name not in seen[kind]:
                plugins[kind].append(PluginParams(name))
                seen[kind].append(name)

        for ob in self.objects:
            note("shape", ob.pluginName)
            note("material", ob.material.pluginName)
            note("texture", ob.texture.pluginName)
        for light in self.lights:
            note("light", light.pluginName)
        return plugins

    def rendererConfig(self):
        """scene.js:114-118."""
        return {"filter": self.filter}


def plugin_masks(tracer_config):
    """Lower a tracerConfig() to the category bit masks of sail_config (bit = category id)."""
    def mask(kind, registry):
        m = 0
        for p in tracer_config[kind]:
            m |= 1 << registry[p.name]
        return m
    return {"shape_mask": mask("shape", SHAPE_PLUGINS), "material_mask": mask("material", MATERIAL_PLUGINS),
            "texture_mask": mask("texture", TEXTURE_PLUGINS), "light_mask": mask("light", LIGHT_PLUGINS)}
