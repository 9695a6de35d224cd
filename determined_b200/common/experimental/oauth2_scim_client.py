from determined_b200.experimental._sdk_more import Oauth2ScimClient  # noqa: F401
