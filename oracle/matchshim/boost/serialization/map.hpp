// oracle/matchshim: empty stand-in for <boost/serialization/map.hpp> (test infrastructure); names come from facade.h
