// oracle/matchshim: empty stand-in for <boost/serialization/base_object.hpp> (test infrastructure); names come from facade.h
